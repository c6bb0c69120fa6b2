"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: detection sharding with no collective on the
data path, and the slab split + all-gather of the mesh grid.  The engine is replaced by a stub
that produces index-valued slabs, so only the partition / gather logic is exercised here; the GPU test
(tests/test_gpu_parity.py::test_grid_sharded_equals_single) runs the same function on the real engine."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_shard_range_partitions_exactly():
    sys.path.insert(0, os.path.join(ROOT, "dsp-slam-rgbd_b200"))
    from dspslam.dist import shard_range, slab_range
    for n in (0, 1, 7, 16, 1024, 1025):
        for w in (1, 2, 3, 4, 8):
            cuts = [shard_range(n, r, w) for r in range(w)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1
            slabs = [slab_range(n, r, w) for r in range(w)]
            assert sum(hi - lo for lo, hi, _ in slabs) == n and len({p for _, _, p in slabs}) == 1


class _StubEngine:
    device = torch.device("cpu")

    def grid_sdf(self, code, G, i0=0, i1=None, coord_mode="reference_shear", out=None, mode=None):
        out.copy_(torch.arange(i0, i1, dtype=torch.float32) * 0.5 + float(code[0]))
        return out


class _StubOptimizer:
    def __init__(self, rank):
        self.rank = rank

    def reconstruct_objects(self, dets):
        from reconstruct.utils import ForceKeyErrorDict
        return [ForceKeyErrorDict(is_good=(d % 5 != 0), t_cam_obj=None if d % 5 == 0 else np.full((4, 4), d, np.float32),
                                  code=None if d % 5 == 0 else np.full(64, self.rank, np.float32), loss=float(d))
                for d in dets]


def _worker(rank, world, port, q):
    for p in (os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dspslam.dist import grid_sdf_sharded, reconstruct_sharded, shard_range
    ok = True
    for G in (4, 5, 9):          # 125 and 729 voxels do not divide by 2: padded slabs
        for pieces in (4, 1, 3):  # one all-gather per piece, overlapped with the next piece's kernel on the GPU box
            vol = grid_sdf_sharded(_StubEngine(), np.array([3.0], np.float32), G, pieces=pieces)
            ok &= vol.shape == (G, G, G)
            ok &= bool(torch.equal(vol.reshape(-1), torch.arange(G ** 3, dtype=torch.float32) * 0.5 + 3.0))
    dets = list(range(11))
    outs = reconstruct_sharded(_StubOptimizer(rank), dets)
    ok &= len(outs) == 11
    for d, o in zip(dets, outs):
        owner = [r for r in range(world) if shard_range(11, r, world)[0] <= d < shard_range(11, r, world)[1]][0]
        ok &= o.loss == float(d) and o.is_good == (d % 5 != 0)
        if o.is_good:
            ok &= float(o.t_cam_obj[0, 0]) == d and float(o.code[0]) == owner
        else:
            ok &= o.t_cam_obj is None and o.code is None
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_world2_gloo_grid_gather_and_detection_sharding():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=150) for _ in procs)
    for p in procs:
        p.join(30)
    assert res == [(0, True), (1, True)]

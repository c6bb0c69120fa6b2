"""Multi-GPU plumbing (one process per GPU, torch.distributed): SURVEY.md 8(e).

* Detections are independent (`src/LocalMapping_util.cc:165-206` loops over them sequentially), so they are
  split contiguously over the ranks and optimised with NO data-path collective; only the results
  (4x4 pose + code + status + loss per object) are gathered on the host side.
* The mesh grid (`reconstruct/optimizer.py:224-228`) is split into contiguous slabs of linear voxel indices -
  slabs of the slowest axis x, the contiguous equivalent of the z-slabs the reference layout would make strided
  (`reconstruct/utils.py:106`: z is the fastest axis) - evaluated with `dspslam_grid_sdf(i0, i1)` on each rank
  and gathered with all-gathers that overlap the remaining slab kernels (NCCL over NVLink on the B200 box; gloo in
  the CPU tests).
"""
from __future__ import annotations

import numpy as np


def shard_range(n, rank, world):
    """Contiguous [lo, hi) of `n` items owned by `rank`: sizes differ by at most one, earlier ranks larger."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def slab_range(total, rank, world):
    """Equal-size slabs (all-gather needs equal counts): [lo, hi) clipped to `total`, and the padded slab size."""
    per = (total + world - 1) // world
    lo = min(rank * per, total)
    return lo, min(lo + per, total), per


def _world(group):
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def piece_ranges(total, rank, world, pieces):
    """The volume is cut into `pieces` contiguous pieces, each piece into `world` equal sub-slabs: rank r evaluates
    sub-slab r of every piece, and ONE all-gather per piece assembles a contiguous range of the final volume.
    Returns (per, [(lo, hi, out_lo)] per piece): `per` voxels per sub-slab (padded), this rank's global index range
    clipped to `total`, and the first index of the piece in the padded output buffer."""
    per = (total + pieces * world - 1) // (pieces * world)
    out = []
    for j in range(pieces):
        lo = min((j * world + rank) * per, total)
        out.append((lo, min(lo + per, total), j * world * per))
    return per, out


def grid_sdf_sharded(engine, code, G, coord_mode="reference_shear", mode=None, group=None, pieces=4, timing=None):
    """(G, G, G) SDF volume on every rank (reconstruct/optimizer.py:224-228 over the ranks).  With W > 1 ranks the
    grid kernel runs piece by piece and the all-gather of piece j (async, NCCL's own stream) overlaps the kernel of
    piece j + 1, so only the last piece's gather is exposed.  `engine` needs `.grid_sdf(code, G, i0, i1, coord_mode=,
    out=, mode=)` and `.device` (dspslam.Engine).  `timing`: optional dict, filled with CUDA-event milliseconds of
    the slab kernels and of the whole call (device timed)."""
    import torch
    import torch.distributed as dist
    rank, world = _world(group)
    total = G ** 3
    if world == 1:
        pieces = 1
    per, ranges = piece_ranges(total, rank, world, pieces)
    full = torch.empty(per * world * pieces, dtype=torch.float32, device=engine.device)
    on_gpu = full.device.type == "cuda"
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)] if (timing is not None and on_gpu) else None
    if ev:
        ev[0].record()
    works = []
    for lo, hi, out_lo in ranges:
        mine = full[out_lo + rank * per: out_lo + (rank + 1) * per]
        if hi > lo:
            engine.grid_sdf(code, G, i0=lo, i1=hi, coord_mode=coord_mode, out=mine[: hi - lo], mode=mode)
        if world > 1:
            works.append(dist.all_gather_into_tensor(full[out_lo: out_lo + world * per], mine if on_gpu else mine.clone(),
                                                     group=group, async_op=True))
    if ev:
        ev[1].record()            # behind the last slab kernel, in front of the waits: the compute part alone
    for w in works:
        w.wait()
    if timing is not None and on_gpu:
        end = torch.cuda.Event(enable_timing=True)
        end.record()
        end.synchronize()
        timing["ms_compute"] = ev[0].elapsed_time(ev[1])
        timing["ms_total"] = ev[0].elapsed_time(end)
        timing["ms_allgather_exposed"] = ev[1].elapsed_time(end)
        timing["pieces"] = pieces
    return full[:total].view(G, G, G)


def reconstruct_sharded(optimizer, detections, group=None):
    """`Optimizer.reconstruct_objects` over this rank's contiguous shard of `detections`; returns the full result
    list on every rank (host-side gather of the small per-object results; no data-path collective)."""
    import torch.distributed as dist
    rank, world = _world(group)
    lo, hi = shard_range(len(detections), rank, world)
    mine = optimizer.reconstruct_objects(detections[lo:hi]) if hi > lo else []
    if world == 1:
        return mine
    packed = [(bool(o.is_good), None if o.t_cam_obj is None else np.asarray(o.t_cam_obj),
               None if o.code is None else np.asarray(o.code), float(o.loss)) for o in mine]
    parts = [None] * world
    dist.all_gather_object(parts, packed, group=group)
    from reconstruct.utils import ForceKeyErrorDict
    return [ForceKeyErrorDict(is_good=g, t_cam_obj=T, code=c, loss=l) for part in parts for (g, T, c, l) in part]

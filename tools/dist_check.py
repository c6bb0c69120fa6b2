"""Multi-GPU check + timing of the mesh-grid path (C4): z/x-slab sharding + NCCL all-gather.
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/dist_check.py [G]
Every rank evaluates its slab with the tcgen05 forward kernel; the slabs are gathered with
all_gather_into_tensor; rank 0 checks the volume against a single-GPU evaluation (bit-exact: rows are
independent) and prints one JSON line with the device-timed phases (max over ranks)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.dist import grid_sdf_sharded, slab_range  # noqa: E402
from dspslam.engine import Engine  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 256
rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = Engine(synth.load_wfit(), device=local, mode="tc")
code = np.zeros(64, np.float32)
code[:3] = [0.2, -0.1, 0.3]
for _ in range(3):
    vol = grid_sdf_sharded(eng, code, G, coord_mode="reference_shear")
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
reps = 5
ev[0].record()
for _ in range(reps):
    vol = grid_sdf_sharded(eng, code, G, coord_mode="reference_shear")
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / reps
# compute-only time of this rank's slab
lo, hi, per = slab_range(G ** 3, rank, world)
buf = torch.empty(per, dtype=torch.float32, device="cuda")
ev[0].record()
for _ in range(reps):
    eng.grid_sdf(code, G, i0=lo, i1=hi, out=buf[: hi - lo])
ev[1].record()
torch.cuda.synchronize()
ms_c = ev[0].elapsed_time(ev[1]) / reps
t = torch.tensor([ms, ms_c], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ref = eng.grid_sdf(code, G, coord_mode="reference_shear").view(G, G, G)
    exact = bool(torch.equal(ref, vol))
    print(json.dumps({"workload": "C4 mesh grid G=%d, slabs of the slowest axis + NCCL all-gather" % G, "n_gpus": world,
                      "ms_total": float(t[0]), "ms_slab_compute": float(t[1]), "ms_gather_and_overhead": float(t[0] - t[1]),
                      "mquery_per_s": G ** 3 / float(t[0]) / 1e3, "tflops": G ** 3 * 3.67104e6 / float(t[0]) / 1e9,
                      "gathered_equals_single_gpu": exact, "has_zero_crossing": bool(vol.min() < 0 < vol.max())}), flush=True)
    assert exact
if world > 1:
    dist.destroy_process_group()

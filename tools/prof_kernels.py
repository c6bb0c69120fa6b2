"""Short single-kernel driver for ncu: one forward-only launch (grid SDF, G=128) and one SDF+Jacobian launch."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402

eng = Engine(synth.load_wfit(), mode="tc")
code = np.zeros(64, np.float32)
code[:3] = [0.2, -0.1, 0.3]
G = int(sys.argv[1]) if len(sys.argv) > 1 else 128
nj = int(sys.argv[2]) if len(sys.argv) > 2 else 148 * 128 * 4
x = torch.rand((nj, 3), device="cuda") * 1.6 - 0.8
for _ in range(2):
    out = eng.grid_sdf(code, G, coord_mode="lattice")
    s, j = eng.sdf_jac(code, x)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
ev[0].record()
out = eng.grid_sdf(code, G, coord_mode="lattice")
ev[1].record()
s, j = eng.sdf_jac(code, x)
ev[2].record()
torch.cuda.synchronize()
t0, t1 = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
print("grid G=%d: %.3f ms  %.1f Mquery/s  %.1f TFLOP/s" % (G, t0, G ** 3 / t0 / 1e3, G ** 3 * 3.67104e6 / t0 / 1e9))
print("sdf_jac n=%d: %.3f ms  %.1f Mquery/s  %.1f TFLOP/s" % (nj, t1, nj / t1 / 1e3, nj * 7.34208e6 / t1 / 1e9))

"""Smallest possible exercise of the tcgen05 path: n rows of sdf / sdf_jac against the oracle (prints, no asserts)."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402
from oracle import dspslam_oracle as orc  # noqa: E402

np.set_printoptions(precision=5, suppress=True, linewidth=160)
dec = synth.load_wfit()
od = orc.OracleDecoder(dec.state_dict())
eng = Engine(dec, mode="tc")
rng = np.random.default_rng(0)
code = (0.3 * rng.normal(size=64)).astype(np.float32)
for n in [int(a) for a in sys.argv[1:]] or [128, 1000, 40000]:
    x = rng.uniform(-0.8, 0.8, size=(n, 3)).astype(np.float32)
    so, jo = od.sdf_jacobian(code, x)
    s = eng.sdf(code, x, mode="tc").cpu().numpy()
    torch.cuda.synchronize()
    print("n=%d sdf: max err %.3e   got %s want %s" % (n, np.abs(s - so).max(), s[:4], so[:4]), flush=True)
    s2, j = eng.sdf_jac(code, x, mode="tc")
    torch.cuda.synchronize()
    s2, j = s2.cpu().numpy(), j.cpu().numpy()
    e = np.abs(j - jo)
    print("n=%d jac: sdf err %.3e  jac max %.3e rms %.3e  rows>1e-3 %.4f" % (n, np.abs(s2 - so).max(), e.max(), np.sqrt((e ** 2).mean()), (e.max(1) > 1e-3).mean()), flush=True)
    print("   row0 got ", j[0, 60:], "\n   row0 want", jo[0, 60:], flush=True)
print("probe done")

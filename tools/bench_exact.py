"""Stand-alone timing of the exact SDF+Jacobian kernel (k_tc_jacx) and, with the `make -C csrc exp` variant libraries,
of its timing experiments (wrong numerics on purpose: 2 no activation stores, 3 no lo(W) pass).
    python tools/bench_exact.py            # runs itself once per library found"""
import os
import subprocess
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
LIBDIR = os.path.join(ROOT, "dsp-slam-rgbd_b200", "dspslam")

if len(sys.argv) > 1:
    from dspslam import _native as N
    N.LIB_PATH = sys.argv[1]
    import numpy as np
    import torch
    from dspslam import synth
    from dspslam.engine import Engine
    eng = Engine(synth.load_wfit())
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    out = [os.path.basename(sys.argv[1])]
    for mode, n in (("tc_exact", 148 * 64 * 4), ("tc", 148 * 128 * 4), ("tc_fast", 148 * 128 * 4)):
        x = torch.rand((n, 3), device="cuda") * 1.6 - 0.8
        for _ in range(2):
            eng.sdf_jac(code, x, mode=mode)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(5):
            eng.sdf_jac(code, x, mode=mode)
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / 5
        out.append("%s n=%d: %.3f ms = %.1f us per tile-trip, %.1f Mquery/s" % (mode, n, ms, ms * 1e3 / 4, n / ms / 1e3))
    print(" | ".join(out), flush=True)
else:
    for lib in ["libdspslam_sm100.so"] + ["libdspslam_sm100_exp%d.so" % v for v in (2, 3)]:
        path = os.path.join(LIBDIR, lib)
        if os.path.isfile(path):
            subprocess.call([sys.executable, os.path.abspath(__file__), path])

"""Noise floor of the reference's own free-running Gauss-Newton trajectory.  TEST INFRASTRUCTURE
(build container only: imports /root/reference).      Usage: python oracle/noise_floor.py

Runs the UNMODIFIED reference `reconstruct_object` on the C1 golden detections three ways that are all
"the reference in FP32": 8 torch threads (the goldens), 1 torch thread (different GEMM blocking, hence
different FP32 rounding), and 8 threads with the surface points perturbed by one FP32 ulp.  The spread of
the final pose / code between these runs is the tolerance floor any other implementation can be held to
for the END state of a free run (the row selections |s|<th, de_do>1e-2 make the trajectory discontinuous
in its inputs - SURVEY.md 8(a) quirk 9).  Writes tests/golden/noise_floor.json.
"""
import contextlib
import io
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

ref = ref_harness.import_reference()
sys.path.append(os.path.join(ROOT, "dsp-slam-rgbd_b200"))
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from make_goldens import ref_decoder, ref_config  # noqa: E402

warnings.simplefilter("ignore")


def run(opt, g, pts):
    with contextlib.redirect_stdout(io.StringIO()):
        out = opt.reconstruct_object(g["in_t_cam_obj"].copy(), pts, g["in_rays"].copy(), g["in_depth"].copy())
    return np.asarray(out.t_cam_obj, np.float32), np.asarray(out.code, np.float32)


def main():
    dec = ref_decoder(synth.load_wfit().state_dict())
    rows = []
    for kind, names in (("kitti", ["c1_obj%d" % k for k in range(4)]),):
        opt = ref.optimizer.Optimizer(dec, ref_config(kind))
        for n in names:
            g = dict(np.load(os.path.join(ROOT, "tests", "golden", n + ".npz")))
            torch.set_num_threads(8)
            Ta, ca = run(opt, g, g["in_pts"].copy())
            torch.set_num_threads(1)
            Tb, cb = run(opt, g, g["in_pts"].copy())
            torch.set_num_threads(8)
            Tc, cc = run(opt, g, np.nextafter(g["in_pts"], np.float32(np.inf)).astype(np.float32))
            for tag, T, c in (("1thread", Tb, cb), ("ulp", Tc, cc)):
                dt, ang, _ = synth.pose_errors(T, Ta)
                rows.append(dict(obj=n, variant=tag, dt_m=float(dt), rot_deg=float(ang),
                                 dcode_rel=float(np.linalg.norm(c - ca) / max(np.linalg.norm(ca), 1e-9)),
                                 vs_golden_rot_deg=float(synth.pose_errors(Ta, g["t_cam_obj"])[1])))
                print(rows[-1], flush=True)
    with open(os.path.join(ROOT, "tests", "golden", "noise_floor.json"), "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()

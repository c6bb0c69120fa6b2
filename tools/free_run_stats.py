"""Free-running end-state agreement between the tensor-core mode and the FP32 validation mode (which follows
the reference to ~1e-6 per step) over N synthetic C1 detections.  Prints the distribution of the differences and
the fraction inside the north_star tolerances (1 cm, 0.1 deg, 1e-2 relative code).
Usage: python tools/free_run_stats.py [N] [iters]"""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    mode_b = sys.argv[3] if len(sys.argv) > 3 else "tc"
    eng = Engine(synth.load_wfit())
    cfg = synth.kitti_config()
    j = cfg.optimizer.joint_optim
    p = dict(k1=j.k1, k2=j.k2, k3=j.k3, k4=j.k4, b1=j.b1, b2=j.b2, lr=j.learning_rate, s_damp=j.scale_damping,
             cut_off=cfg.optimizer.cut_off_threshold, num_iterations=iters,
             num_depth_samples=cfg.optimizer.num_depth_samples)
    dets = [synth.make_kitti_detection(k) for k in range(n)]
    a = eng.gn_joint(dets, p, mode="fp32")
    b = eng.gn_joint(dets, p, mode=mode_b)
    dt, ang, dc = [], [], []
    for k in range(n):
        if a.status[k] or b.status[k]:
            continue
        t, r, _ = synth.pose_errors(b.t_cam_obj[k], a.t_cam_obj[k])
        dt.append(t)
        ang.append(r)
        dc.append(np.linalg.norm(b.code[k] - a.code[k]) / max(1e-12, np.linalg.norm(a.code[k])))
    dt, ang, dc = map(np.array, (dt, ang, dc))
    ok = (dt < 1e-2) & (ang < 0.1) & (dc < 1e-2)
    q = lambda v: "median %.2e  p90 %.2e  max %.2e" % (np.median(v), np.quantile(v, 0.9), v.max())
    print("[%s vs fp32] objects compared: %d (status ok in both modes), %d GN iterations" % (mode_b, len(dt), iters))
    print("|dt| [m]   ", q(dt))
    print("rot [deg]  ", q(ang))
    print("code rel   ", q(dc))
    print("within 1 cm / 0.1 deg / 1e-2: %.3f   (pose only: %.3f)" % (ok.mean(), ((dt < 1e-2) & (ang < 0.1)).mean()))
    print("loss: fp32 mean %.5f  tc mean %.5f  max |diff| %.2e" % (a.loss.mean(), b.loss.mean(), np.abs(a.loss - b.loss).max()))


if __name__ == "__main__":
    main()

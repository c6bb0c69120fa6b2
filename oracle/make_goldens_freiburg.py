"""Extra golden vectors: `reconstruct_object` with the reference's Freiburg (monocular cars) hyper-parameters
(`configs/config_freiburg_001.json`: k1 = 1, k3 = 0.5, k4 = 0, scale_damping = 100) - the third dataset
configuration the reference ships, next to KITTI (c1_*) and Redwood (c3_*).  TEST INFRASTRUCTURE, build container
only; reuses the tracing harness of make_goldens.py.      Usage: python oracle/make_goldens_freiburg.py
-> tests/golden/c3f_obj<k>.npz  (the reference's own 50 depth samples, 10 iterations, object scale ~2 at 6-12 m)"""
import os

import numpy as np

import make_goldens as mg          # imports the unmodified reference
from make_goldens import ref, synth, torch


def main():
    with np.load(synth.WFIT_PATH) as z:
        wfit = {k: torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
    dec = mg.ref_decoder(wfit)
    cfg = ref.utils.get_configs(os.path.join(mg.ref_harness.REF_ROOT, "configs/config_freiburg_001.json"))
    cfg.optimizer.joint_optim.num_iterations = 10
    opt = ref.optimizer.Optimizer(dec, cfg)
    for k in range(2):
        det = synth.make_mono_detection(100 + k, M=128, n_fg=300, n_bg=100, scale_range=(1.8, 2.2),
                                        dist_range=(6.0, 12.0))
        g = mg.run_recon(opt, det, det["code"])
        np.savez_compressed(os.path.join(mg.OUT, "c3f_obj%d.npz" % k), **g)
        if g["is_good"]:
            dt, ang, sr = synth.pose_errors(g["t_cam_obj"], det["t_cam_obj_gt"])
            print("c3f obj", k, "K", g["K"].tolist(), "loss %.5f" % g["loss"],
                  "vs GT: %.3f m %.2f deg scale x%.3f" % (dt, ang, sr))
        else:
            print("c3f obj", k, "FAILED", g["K"].tolist())


if __name__ == "__main__":
    main()

"""Generate the LARGE golden sets by running the UNMODIFIED reference (`/root/reference`, torch CPU, FP32):

  bulk_c1.npz      128 C1 detections (KITTI car, M=250, N=450, 10 iterations; seeds 1234 + k, k = 0..127):
                   per-iteration reference state (T_oc, code), b, K, final pose / code / loss.  Inputs are NOT
                   stored: `dspslam.synth.make_kitti_detection(k)` regenerates them; `in_sum[k]` (float64 sum of
                   all inputs) guards against a drifting generator.  H is kept for the first 8 only.
  bulk_c3.npz      32 mono (Redwood hyper-parameters, 48 samples, 20 iterations) detections, same content.
  c2_obj<k>.npz    4 BASELINE-C2-sized detections (M=2048, N=2248, D=50): full per-iteration H, b, K, state.
  grid_sdf_g64.npz MeshExtractor grid SDF at G=64 (reference shear) for the optimised code of C1 object 0.

The population is what the ">= 95 % of reference objects within tolerance" statistic of north_star is taken over
(tests/test_gpu_parity_bulk.py, bench.py `parity`).
TEST INFRASTRUCTURE; runs only in the build container.   Usage: python oracle/make_goldens_bulk.py [n_c1] [n_c3]"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_goldens as mg  # noqa: E402  (imports the reference through ref_harness)

ref, synth, torch = mg.ref, mg.synth, mg.torch
OUT = mg.OUT


def in_sum(det):
    return float(sum(np.asarray(det[k], np.float64).sum() for k in ("t_cam_obj", "pts", "rays", "depth")))


def bulk(opt, make, n, n_it, with_code, keep_H=8):
    keys = dict(T_oc=[], code_it=[], b=[], K=[], t_cam_obj=[], code=[], loss=[], is_good=[], in_sum=[], H=[])
    for k in range(n):
        det = make(k)
        g = mg.run_recon(opt, det, det["code"] if with_code else None)
        good = bool(g["is_good"]) and g["T_oc"].shape[0] == n_it
        keys["is_good"].append(good)
        keys["in_sum"].append(in_sum(det))
        pad = lambda a, shp: a if good else np.full(shp, np.nan, np.float32)
        keys["T_oc"].append(pad(g["T_oc"], (n_it, 4, 4)))
        keys["code_it"].append(pad(g["code_it"], (n_it, 64)))
        keys["b"].append(pad(g["b"], (n_it, 71)))
        keys["K"].append(g["K"] if good else np.full(n_it, -1))
        keys["t_cam_obj"].append(g["t_cam_obj"] if good else np.full((4, 4), np.nan, np.float32))
        keys["code"].append(g["code"] if good else np.full(64, np.nan, np.float32))
        keys["loss"].append(float(g["loss"]))
        if k < keep_H:
            keys["H"].append(pad(g["H"], (n_it, 71, 71)))
        if k % 8 == 0:
            print("  obj", k, "good", good, "K", g["K"].tolist()[:4], "...", flush=True)
    return {k: np.array(v) for k, v in keys.items()}


def main():
    n_c1 = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    n_c3 = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    with np.load(synth.WFIT_PATH) as z:
        wfit = {k: torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
    dec = mg.ref_decoder(wfit)

    opt = ref.optimizer.Optimizer(dec, mg.ref_config("kitti"))
    # ---- C2-sized detections ------------------------------------------------------------------------
    for k in range(4):
        det = synth.make_kitti_detection(k, M=2048, n_bg=200)
        g = mg.run_recon(opt, det)
        for drop in ("in_pts", "in_rays", "in_depth", "in_t_cam_obj"):     # regenerated from the seed
            g.pop(drop)
        g["in_sum"] = np.array(in_sum(det))
        np.savez_compressed(os.path.join(OUT, "c2_obj%d.npz" % k), **g)
        print("c2 obj", k, "good", bool(g["is_good"]), "K", g["K"].tolist(), flush=True)

    # ---- grid at G = 64 -----------------------------------------------------------------------------
    g0 = mg.run_recon(opt, synth.make_kitti_detection(0))
    code = g0["code"].astype(np.float32)
    me = ref.optimizer.MeshExtractor(dec, 64, 64)
    sdf = ref.loss_utils.decode_sdf(dec, torch.from_numpy(code), me.voxel_points)
    np.savez_compressed(os.path.join(OUT, "grid_sdf_g64.npz"), code=code, G=np.array(64),
                        sdf=sdf.numpy().reshape(64, 64, 64))
    print("grid G=64 done: sdf range", float(sdf.min()), float(sdf.max()), flush=True)

    # ---- bulk populations ---------------------------------------------------------------------------
    print("bulk C1 x", n_c1, flush=True)
    np.savez_compressed(os.path.join(OUT, "bulk_c1.npz"),
                        **bulk(opt, lambda k: synth.make_kitti_detection(k), n_c1, 10, False))
    opt3 = ref.optimizer.Optimizer(dec, mg.ref_config("redwood"))
    print("bulk C3 x", n_c3, flush=True)
    np.savez_compressed(os.path.join(OUT, "bulk_c3.npz"),
                        **bulk(opt3, lambda k: synth.make_mono_detection(k), n_c3, 20, True))
    print("done")


if __name__ == "__main__":
    main()

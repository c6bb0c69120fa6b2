"""GPU diagnostics: run every stage of the CUDA path against the goldens / oracle and PRINT the
errors (no asserts), so one gpurun call shows where a problem is.  Usage: python tools/gpu_diag.py [fp32|tc ...]"""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402
from oracle import dspslam_oracle as orc  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
np.set_printoptions(precision=4, suppress=True, linewidth=160)


def load(n):
    with np.load(os.path.join(G, n)) as z:
        return {k: z[k] for k in z.files}


def params(cfg):
    j = cfg.optimizer.joint_optim
    return dict(k1=j.k1, k2=j.k2, k3=j.k3, k4=j.k4, b1=j.b1, b2=j.b2, lr=j.learning_rate, s_damp=j.scale_damping,
                cut_off=cfg.optimizer.cut_off_threshold, num_iterations=j.num_iterations,
                num_depth_samples=cfg.optimizer.num_depth_samples)


def section(name):
    print("\n=== %s ===" % name, flush=True)


def guarded(fn, *a):
    try:
        fn(*a)
    except Exception:
        traceback.print_exc()
    torch.cuda.synchronize()


def diag_queries(eng, mode, which, dec_sd):
    g = load("queries_%s.npz" % which)
    sdf = eng.sdf(g["code"], g["xyz"], mode=mode).cpu().numpy()
    s2, jac = eng.sdf_jac(g["code"], g["xyz"], mode=mode)
    s2, jac = s2.cpu().numpy(), jac.cpu().numpy()
    ej = np.abs(jac - g["jac"])
    print("[%s/%s] sdf max err %.3e | sdf(jac kernel) %.3e | jac max %.3e rms %.3e | rows with entry err>1e-3: %.4f"
          % (mode, which, np.abs(sdf - g["sdf"]).max(), np.abs(s2 - g["sdf_j"]).max(), ej.max(),
             np.sqrt((ej ** 2).mean()), (ej.max(1) > 1e-3).mean()))
    if ej.max() > 1e-2:
        r = np.argmax(ej.max(1))
        print("   worst row", r, "got", jac[r, 60:], "want", g["jac"][r, 60:])
    # larger ragged n (tail handling)
    rng = np.random.default_rng(3)
    od = orc.OracleDecoder(dec_sd)
    for n in (1, 31, 33, 129, 1000):
        x = rng.uniform(-0.8, 0.8, size=(n, 3)).astype(np.float32)
        a = eng.sdf(g["code"], x, mode=mode).cpu().numpy()
        b, jb = eng.sdf_jac(g["code"], x, mode=mode)
        so, jo = od.sdf_jacobian(g["code"], x)
        print("   n=%4d sdf err %.2e / %.2e  jac err %.2e" % (n, np.abs(a - so).max(), np.abs(b.cpu().numpy() - so).max(),
                                                              np.abs(jb.cpu().numpy() - jo).max()))


def diag_grid(eng, mode, dec_sd):
    g = load("grid_sdf.npz")
    Gd = int(g["G"])
    out = eng.grid_sdf(g["code"], Gd, mode=mode).cpu().numpy().reshape(Gd, Gd, Gd)
    print("[%s] grid G=%d shear: max err %.3e" % (mode, Gd, np.abs(out - g["sdf"]).max()))
    od = orc.OracleDecoder(dec_sd)
    lat = orc.grid_sdf(od, g["code"], 12, "lattice")
    out = eng.grid_sdf(g["code"], 12, coord_mode="lattice", mode=mode).cpu().numpy().reshape(12, 12, 12)
    print("[%s] grid G=12 lattice: max err %.3e" % (mode, np.abs(out - lat).max()))
    part = eng.grid_sdf(g["code"], Gd, i0=1000, i1=1777, mode=mode).cpu().numpy()
    print("[%s] grid slab [1000,1777): max err %.3e" % (mode, np.abs(part - g["sdf"].reshape(-1)[1000:1777]).max()))


def diag_step(eng, mode, name, cfg, steps):
    g = load(name)
    p = params(cfg)
    p["num_iterations"] = 1
    for e in steps:
        T_co = np.linalg.inv(g["T_oc"][e].astype(np.float64)).astype(np.float32)
        res = eng.gn_joint([(T_co, g["in_pts"], g["in_rays"], g["in_depth"], g["code_it"][e])], p, mode=mode, debug=True)
        d = res.dbg[0, 0]
        H, b = d[:5041].reshape(71, 71), d[5041:5112]
        n_it = g["H"].shape[0]
        if e + 1 < n_it:
            Tn, cn = np.linalg.inv(g["T_oc"][e + 1].astype(np.float64)), g["code_it"][e + 1]
        else:
            Tn, cn = g["t_cam_obj"], g["code"]
        dt, ang, _ = synth.pose_errors(res.t_cam_obj[0], Tn)
        print("[%s] %s it %2d: status %d K %d (ref %d) n_valid %d | H rel %.2e  b rel %.2e | next pose dt %.2e m rot %.2e deg  dcode %.2e | loss %.5f"
              % (mode, name, e, res.status[0], int(d[5112]), g["K"][e], int(d[5113]),
                 np.linalg.norm(H - g["H"][e]) / np.linalg.norm(g["H"][e]),
                 np.linalg.norm(b - g["b"][e]) / np.linalg.norm(g["b"][e]), dt, ang,
                 np.linalg.norm(res.code[0] - cn), res.loss[0]))
        if np.linalg.norm(H - g["H"][e]) / np.linalg.norm(g["H"][e]) > 1e-2 and e == steps[0]:
            print("  H[:7,:7] got\n", H[:7, :7], "\n  want\n", g["H"][e][:7, :7])
            print("  b[:10] got", b[:10], "\n  want  ", g["b"][e][:10])
            print("  sdf_loss %.6f render_loss %.6f scale %.5f" % (d[5114], d[5115], d[5116]))


def diag_free(eng, mode, cfg, names, with_code=False):
    p = params(cfg)
    gs = [load(n) for n in names]
    dets = [(g["in_t_cam_obj"], g["in_pts"], g["in_rays"], g["in_depth"], g["in_code"] if with_code else None) for g in gs]
    torch.cuda.synchronize()
    t0 = time.time()
    res = eng.gn_joint(dets, p, mode=mode)
    t1 = time.time()
    for i, g in enumerate(gs):
        if res.status[i] == 0 and bool(g["is_good"]):
            dt, ang, _ = synth.pose_errors(res.t_cam_obj[i], g["t_cam_obj"])
            print("[%s] free run %s: dt %.2e m rot %.3f deg  |dcode| %.3e (|code| %.3f)  loss %.5f (ref %.5f)"
                  % (mode, names[i], dt, ang, np.linalg.norm(res.code[i] - g["code"]), np.linalg.norm(g["code"]),
                     res.loss[i], g["loss"]))
        else:
            print("[%s] free run %s: status %d (ref is_good %s)" % (mode, names[i], res.status[i], g["is_good"]))
    print("   batch of %d took %.1f ms (incl. H2D/D2H)" % (len(gs), 1e3 * (t1 - t0)))


def diag_pose_only(eng, mode):
    for k in range(2):
        g = load("pose_only_obj%d.npz" % k)
        T, dbg = eng.gn_pose_only([(g["in_t_co_se3"], float(g["in_scale"]), g["in_pts"], g["in_code"])], 5, mode=mode,
                                  debug=True)
        H0 = dbg[0, 0, :36].reshape(6, 6)
        b0 = dbg[0, 0, 36:42]
        print("[%s] pose-only %d: H0 rel %.2e b0 rel %.2e | final T max abs diff %.2e"
              % (mode, k, np.linalg.norm(H0 - g["H"][0]) / np.linalg.norm(g["H"][0]),
                 np.linalg.norm(b0 - g["b"][0]) / np.linalg.norm(g["b"][0]), np.abs(T[0] - g["t_cam_obj"]).max()))


def main():
    modes = sys.argv[1:] or ["fp32"]
    print(torch.cuda.get_device_name(0))
    dec = synth.load_wfit()
    dec_r = synth.make_decoder(0)
    eng = Engine(dec, mode="fp32")
    eng_r = Engine(dec_r, mode="fp32")
    for mode in modes:
        section("queries " + mode)
        guarded(diag_queries, eng, mode, "wfit", dec.state_dict())
        guarded(diag_queries, eng_r, mode, "wrand", dec_r.state_dict())
        section("grid " + mode)
        guarded(diag_grid, eng, mode, dec.state_dict())
        section("teacher-forced GN steps " + mode)
        guarded(diag_step, eng, mode, "c1_obj0.npz", synth.kitti_config(), [0, 1, 5, 9])
        guarded(diag_step, eng, mode, "c1_obj2.npz", synth.kitti_config(), [0, 9])
        guarded(diag_step, eng, mode, "c3_obj0.npz", synth.redwood_config(), [0, 10, 19])
        section("free runs " + mode)
        guarded(diag_free, eng, mode, synth.kitti_config(), ["c1_obj%d.npz" % k for k in range(4)])
        guarded(diag_free, eng, mode, synth.redwood_config(), ["c3_obj%d.npz" % k for k in range(2)], True)
        guarded(diag_free, eng_r, mode, synth.kitti_config(), ["wrand_fail.npz"])
        section("pose only " + mode)
        guarded(diag_pose_only, eng, mode)
        g = load("objectpoint.npz")
        guarded(lambda: print("[%s] objectpoint mean sdf %.7f (ref %.7f)" % (mode, eng.mean_sdf(g["code"], g["pts"], mode=mode).item(), g["mean_sdf"])))
    print("launches:", eng.launches)


if __name__ == "__main__":
    main()

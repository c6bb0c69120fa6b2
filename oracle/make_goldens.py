"""Generate tests/golden/*.npz by running the UNMODIFIED reference code
(`/root/reference`, torch CPU, FP32) on seeded synthetic inputs.
TEST INFRASTRUCTURE; runs only in the build container (the reference does not
travel to the GPU box, the vectors do).      Usage: python oracle/make_goldens.py

What is captured (all FP32 unless noted):
  queries_<w>.npz      decode_sdf + get_batch_sdf_jacobian at random points (W-rand, W-fit)
  liealg.npz           exp_sim3 / exp_se3 / huber / rotation-prior known answers
  voxel_grid.npz       create_voxel_grid for several G (sheared grid of modern torch)
  grid_sdf.npz         MeshExtractor grid SDF (G=16) for one code
  c1_obj<k>.npz        reconstruct_object on KITTI-style cars: per-iteration state (T_oc, code), H, b, K
  c3_obj<k>.npz        same with the Redwood (mono) hyper-parameters, 48 depth samples, warm-start code
  pose_only_obj<k>.npz estimate_pose_cam_obj
  wrand_fail.npz       reconstruct_object with W-rand (no zero crossing) -> is_good False
  objectpoint.npz      compute_sdf_loss_objectpoint_zhjd
"""
import contextlib
import io
import os
import sys
import tempfile
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

ref = ref_harness.import_reference()          # puts /root/reference first on sys.path
sys.path.append(os.path.join(ROOT, "dsp-slam-rgbd_b200"))
import torch  # noqa: E402
from dspslam import synth  # noqa: E402  (numpy generators; `reconstruct`/`deep_sdf` resolve to the reference)

OUT = os.path.join(ROOT, "tests", "golden")
warnings.simplefilter("ignore")
torch.set_num_threads(8)


def ref_decoder(state_dict):
    """Load through the reference's own config_decoder (workspace.py:202-223)."""
    tmp = tempfile.mkdtemp()
    dec = ref.decoder.Decoder(synth.CODE_LEN, **synth.NETWORK_SPECS)
    dec.load_state_dict(state_dict)
    synth.write_experiment_dir(dec, tmp)
    cfg = ref.utils.ForceKeyErrorDict(DeepSDF_DIR=tmp)
    d = ref.utils.get_decoder(cfg)
    for p in d.parameters():
        p.requires_grad_(False)
    return d


def ref_config(kind):
    if kind == "kitti":
        return ref.utils.get_configs(os.path.join(ref_harness.REF_ROOT, "configs/config_kitti.json"))
    cfg = ref.utils.get_configs(os.path.join(ref_harness.REF_ROOT, "configs/config_redwood_01053.json"))
    cfg.optimizer.num_depth_samples = 48
    cfg.optimizer.joint_optim.num_iterations = 20
    return cfg


class Trace:
    """Record per-iteration H, b and the render row count while the reference runs."""

    def __enter__(self):
        self.H, self.b, self.K, self.dp, self.T_oc, self.code = [], [], [], [], [], []
        self._sdf = ref.optimizer.compute_sdf_loss
        self._inv, self._mv = torch.inverse, torch.mv
        self._render, self._exp = ref.optimizer.compute_render_loss, ref.optimizer.exp_sim3
        self._exp_se3 = ref.optimizer.exp_se3

        def inv(m):
            if m.shape[0] > 4:
                self.H.append(m.detach().clone().numpy())
            return self._inv(m)

        def mv(a, v):
            if a.shape[0] > 4:
                self.b.append(v.detach().clone().numpy())
            return self._mv(a, v)

        def render(*a, **k):
            r = self._render(*a, **k)
            self.K.append(-1 if r is None else int(r[2].shape[0]))
            return r

        def sdf(dec, pts, t_obj_cam, lat):
            self.T_oc.append(t_obj_cam.detach().clone().numpy())
            self.code.append(lat.detach().clone().numpy())
            return self._sdf(dec, pts, t_obj_cam, lat)

        def exp(x):
            self.dp.append(x.detach().clone().numpy())
            return self._exp(x)

        def exp6(x):
            self.dp.append(x.detach().clone().numpy())
            return self._exp_se3(x)

        torch.inverse, torch.mv = inv, mv
        ref.optimizer.compute_render_loss = render
        ref.optimizer.compute_sdf_loss = sdf
        ref.optimizer.exp_sim3 = exp
        ref.optimizer.exp_se3 = exp6
        return self

    def __exit__(self, *a):
        torch.inverse, torch.mv = self._inv, self._mv
        ref.optimizer.compute_render_loss = self._render
        ref.optimizer.compute_sdf_loss = self._sdf
        ref.optimizer.exp_sim3 = self._exp
        ref.optimizer.exp_se3 = self._exp_se3


def run_recon(opt, det, code=None):
    with Trace() as tr, contextlib.redirect_stdout(io.StringIO()):
        out = opt.reconstruct_object(det["t_cam_obj"].copy(), det["pts"], det["rays"], det["depth"],
                                     code)
    g = dict(in_t_cam_obj=det["t_cam_obj"], in_pts=det["pts"], in_rays=det["rays"],
             in_depth=det["depth"], gt_t_cam_obj=det["t_cam_obj_gt"], gt_code=det["code_gt"],
             H=np.array(tr.H, np.float32).reshape(-1, 71, 71),
             b=np.array(tr.b, np.float32).reshape(-1, 71),
             T_oc=np.array(tr.T_oc, np.float32), code_it=np.array(tr.code, np.float32),
             K=np.array(tr.K, np.int64), dp=np.array(tr.dp, np.float32).reshape(-1, 7),
             is_good=np.array(bool(out.is_good)), loss=np.array(float(out.loss), np.float64))
    if code is not None:
        g["in_code"] = code
    if out.is_good:
        g["t_cam_obj"] = out.t_cam_obj
        g["code"] = out.code
    return g


def main():
    os.makedirs(OUT, exist_ok=True)
    wrand = synth.make_decoder(0).state_dict()
    with np.load(synth.WFIT_PATH) as z:
        wfit = {k: torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
    decs = {"wrand": ref_decoder(wrand), "wfit": ref_decoder(wfit)}

    # ---- per-query SDF / Jacobian -----------------------------------------
    for name, dec in decs.items():
        rng = np.random.default_rng(11)
        n = 600
        xyz = rng.uniform(-0.9, 0.9, size=(n, 3)).astype(np.float32)
        # a third of the points hugging the W-fit surface (where GN evaluates)
        u = rng.normal(size=(n // 3, 3))
        u /= np.linalg.norm(u, axis=1, keepdims=True)
        code = (0.3 * rng.normal(size=synth.CODE_LEN)).astype(np.float32)
        axes, centres = synth.shape_params(code)
        prim = rng.integers(0, 2, size=n // 3)
        xyz[: n // 3] = (centres[prim] + u * axes[prim] * (1 + 0.02 * rng.normal(size=(n // 3, 1)))
                         ).astype(np.float32)
        lat = torch.from_numpy(code)
        sdf = ref.loss_utils.decode_sdf(dec, lat, torch.from_numpy(xyz)).numpy()
        y, jac = ref.loss_utils.get_batch_sdf_jacobian(dec, lat, torch.from_numpy(xyz), 1)
        np.savez(os.path.join(OUT, "queries_%s.npz" % name), code=code, xyz=xyz, sdf=sdf,
                 sdf_j=y.numpy().reshape(-1), jac=jac.numpy().reshape(n, 67))
        print("queries", name, "sdf range", sdf.min(), sdf.max())

    # ---- Lie algebra / robust kernel / rotation prior ------------------------
    rng = np.random.default_rng(5)
    xs = [np.zeros(7), np.array([0.1, -0.2, 0.3, 0, 0, 0, 0.05]), np.array([0.1, -0.2, 0.3, 0, 0, 0, -0.05]),
          np.array([0.3, 0.1, -0.2, 0.02, -0.4, 0.1, 0.2]), np.array([0.3, 0.1, -0.2, 0.02, -0.4, 0.1, -0.2]),
          np.array([1e-3, 2e-3, -1e-3, 1e-4, -2e-4, 3e-4, 1e-3])]
    xs += [rng.normal(scale=0.3, size=7) for _ in range(6)]
    xs = np.array(xs, np.float32)
    sim3 = np.array([ref.loss_utils.exp_sim3(torch.from_numpy(x)).numpy() for x in xs])
    se3 = np.array([ref.loss_utils.exp_se3(torch.from_numpy(x[:6])).numpy() for x in xs])
    r = np.concatenate([[0, 1e-4, -1e-4, 0.02, -0.02, 0.025, 0.3, -0.3],
                        rng.normal(scale=0.05, size=40)]).astype(np.float32)
    hub = {}
    for bb in (0.025, 0.2, 0.05):
        rr, loss, w = ref.loss_utils.get_robust_res(torch.from_numpy(r.copy()), bb)
        hub["hub_r_%g" % bb] = rr.numpy().reshape(-1)
        hub["hub_loss_%g" % bb] = np.array(loss.item(), np.float32)
        hub["hub_w_%g" % bb] = w.numpy().reshape(-1)
    rot_T, rot_J, rot_res = [], [], []
    for k in range(6):
        det = synth.make_kitti_detection(100 + k, M=16, n_bg=0)
        T = det["t_cam_obj"].copy()
        if k >= 3:   # tilt it so the prior is active
            tilt = ref.loss_utils.exp_se3(torch.tensor([0, 0, 0, 0.1 * k, 0.05, -0.07 * k])).numpy()[:3, :3]
            T[:3, :3] = tilt @ T[:3, :3]
        T_oc = torch.inverse(torch.from_numpy(T))
        J, res = ref.loss.compute_rotation_loss_sim3(T_oc.clone())
        rot_T.append(T_oc.numpy())
        rot_J.append(J.numpy())
        rot_res.append(float(res))
    np.savez(os.path.join(OUT, "liealg.npz"), x=xs, exp_sim3=sim3, exp_se3=se3, hub_in=r,
             rot_T_oc=np.array(rot_T), rot_J=np.array(rot_J), rot_res=np.array(rot_res, np.float32), **hub)

    # ---- voxel grid + grid sdf -------------------------------------------------
    vg = {}
    for G in (4, 8, 12, 20, 32):
        vg["G%d" % G] = ref.utils.create_voxel_grid(G).numpy()
    np.savez_compressed(os.path.join(OUT, "voxel_grid.npz"), **vg)
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    me = ref.optimizer.MeshExtractor(decs["wfit"], 64, 16)
    sdf_t = ref.loss_utils.decode_sdf(decs["wfit"], torch.from_numpy(code), me.voxel_points)
    np.savez_compressed(os.path.join(OUT, "grid_sdf.npz"), code=code, G=np.array(16),
                        sdf=sdf_t.numpy().reshape(16, 16, 16))

    # ---- C1: KITTI cars -------------------------------------------------------------
    opt = ref.optimizer.Optimizer(decs["wfit"], ref_config("kitti"))
    for k in range(4):
        det = synth.make_kitti_detection(k)
        g = run_recon(opt, det)
        np.savez_compressed(os.path.join(OUT, "c1_obj%d.npz" % k), **g)
        if g["is_good"]:
            dt, ang, sr = synth.pose_errors(g["t_cam_obj"], det["t_cam_obj_gt"])
            print("c1 obj", k, "K", g["K"].tolist(), "loss %.5f" % g["loss"],
                  "vs GT: %.3f m %.2f deg scale x%.3f" % (dt, ang, sr))
        else:
            print("c1 obj", k, "FAILED", g["K"].tolist())

    # ---- C3: mono (Redwood hyper-parameters) --------------------------------------------
    opt3 = ref.optimizer.Optimizer(decs["wfit"], ref_config("redwood"))
    for k in range(2):
        det = synth.make_mono_detection(k, M=128, n_fg=300, n_bg=100)
        g = run_recon(opt3, det, det["code"])
        np.savez_compressed(os.path.join(OUT, "c3_obj%d.npz" % k), **g)
        if g["is_good"]:
            dt, ang, sr = synth.pose_errors(g["t_cam_obj"], det["t_cam_obj_gt"])
            print("c3 obj", k, "K", g["K"].tolist(), "loss %.5f" % g["loss"],
                  "vs GT: %.3f m %.2f deg scale x%.3f" % (dt, ang, sr))
        else:
            print("c3 obj", k, "FAILED", g["K"].tolist())

    # ---- pose only ----------------------------------------------------------------------
    for k in range(2):
        det = synth.make_kitti_detection(k)
        T0 = det["t_cam_obj"].astype(np.float64)
        s0 = np.cbrt(np.linalg.det(T0[:3, :3]))
        se3 = T0.copy()
        se3[:3, :3] /= s0
        se3 = se3.astype(np.float32)
        code = det["code_gt"]
        with Trace() as tr, contextlib.redirect_stdout(io.StringIO()):
            out = opt.estimate_pose_cam_obj(se3.copy(), float(s0), det["pts"], code)
        np.savez_compressed(os.path.join(OUT, "pose_only_obj%d.npz" % k), in_t_co_se3=se3,
                            in_scale=np.array(s0, np.float64), in_pts=det["pts"], in_code=code,
                            H=np.array(tr.H, np.float32), b=np.array(tr.b, np.float32),
                            T_oc=np.array(tr.T_oc, np.float32),
                            t_cam_obj=out.numpy(), gt_t_cam_obj=det["t_cam_obj_gt"])
        print("pose-only", k, synth.pose_errors(
            np.vstack([out.numpy()[:3] * np.array([[s0, s0, s0, 1]]), [[0, 0, 0, 1]]]), det["t_cam_obj_gt"]))

    # ---- failure path: W-rand has no surface -----------------------------------------------
    optr = ref.optimizer.Optimizer(decs["wrand"], ref_config("kitti"))
    det = synth.make_kitti_detection(0)
    g = run_recon(optr, det)
    print("wrand is_good", g["is_good"], "K", g["K"].tolist())
    np.savez_compressed(os.path.join(OUT, "wrand_fail.npz"), **g)

    # ---- object-point SDF loss -------------------------------------------------------------
    rng = np.random.default_rng(3)
    pts = rng.uniform(-0.6, 0.6, size=(300, 3)).astype(np.float32)
    code = (0.2 * rng.normal(size=64)).astype(np.float32)
    with contextlib.redirect_stdout(io.StringIO()):
        v = opt.compute_sdf_loss_objectpoint_zhjd(pts, code)
    np.savez(os.path.join(OUT, "objectpoint.npz"), pts=pts, code=code, mean_sdf=np.array(float(v), np.float32))
    print("done ->", OUT)


if __name__ == "__main__":
    main()

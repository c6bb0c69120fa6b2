"""Synthetic decoders and detections for parity tests and benchmarks.

The reference ships no weights, no data and no tests, so every
input used to compare against it is generated here, seeded, following
SURVEY.md §8(d):

* ``W-rand``: the seeded default initialisation of the 8x512 decoder.  Its SDF
  has no zero crossing, so it pins per-query values and the failure path only.
* ``W-fit`` : W-rand fitted with Adam to a family of car-like two-ellipsoid
  pseudo-SDFs (so Gauss-Newton converges).  Stored as fp16-rounded
  parameters in ``assets/wfit_decoder_fp16.npz`` because training is not
  bit-reproducible across machines (made by ``tools/make_wfit.py``).
* detections C1/C2/C3: KITTI-style cars and mono objects whose surface points
  and rays come from the same shape family.
"""
from __future__ import annotations

import os

import numpy as np

CODE_LEN = 64
NETWORK_SPECS = dict(
    dims=[512] * 8,
    dropout=list(range(8)),
    dropout_prob=0.2,
    norm_layers=list(range(8)),
    latent_in=[4],
    xyz_in_all=False,
    use_tanh=False,
    latent_dropout=False,
    weight_norm=True,
)
_REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
WFIT_PATH = os.path.join(_REPO, "assets", "wfit_decoder_fp16.npz")

# Training shape family: a car-like union of two ellipsoids (body + cabin) in the
# DeepSDF canonical frame (y up, z along the length, half-length ~0.9).  Two
# primitives make all six pose DOF observable (a single ellipsoid is nearly
# symmetric about its long axis, which makes the reference's GN oscillate).
PRIM_AXES = np.array([[0.45, 0.24, 0.90], [0.38, 0.22, 0.42]], dtype=np.float64)
PRIM_CENTRES = np.array([[0.0, -0.10, 0.0], [0.0, 0.16, -0.12]], dtype=np.float64)


# --------------------------------------------------------------------------
# decoders
# --------------------------------------------------------------------------
def make_decoder(seed=0):
    """W-rand: seeded default init (CPU, eval mode)."""
    import torch
    import warnings
    from deep_sdf.deep_sdf_decoder import Decoder
    torch.manual_seed(seed)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        dec = Decoder(CODE_LEN, **NETWORK_SPECS)
    return dec.eval()


def shape_params(code):
    """(axes (...,2,3), centres (...,2,3)) of the two primitives for a latent code:
    the first three code entries stretch the shape by up to +-10 % per axis."""
    code = np.asarray(code, dtype=np.float64)
    stretch = 1.0 + 0.1 * np.tanh(code[..., None, :3])
    return PRIM_AXES * stretch, PRIM_CENTRES * stretch


def _prim_f_and_grad(x, axes, centres):
    """f = ||(x-c)/a|| - 1 and its gradient for each primitive; x (..., 3)."""
    d = x[..., None, :] - centres
    q = d / axes
    r = np.maximum(np.linalg.norm(q, axis=-1), 1e-9)
    f = r - 1.0
    grad = q / axes / r[..., None]
    return f, grad


def shape_sdf(x, axes, centres, clamp=0.1):
    """First-order distance min_k f_k / ||grad f_k|| of the union, clamped (the
    pseudo-SDF the W-fit decoder is trained on; unit gradient near the surface)."""
    f, g = _prim_f_and_grad(x, axes, centres)
    d = f / np.maximum(np.linalg.norm(g, axis=-1), 1e-9)
    return np.clip(d.min(axis=-1), -clamp, clamp)


def state_dict_to_npz(state_dict, path):
    arrs = {k: v.detach().cpu().numpy().astype(np.float16) for k, v in state_dict.items()}
    np.savez_compressed(path, **arrs)


def load_decoder_npz(path=WFIT_PATH):
    """Decoder whose parameters are the (fp16-representable) values in `path`."""
    import torch
    dec = make_decoder(0)
    with np.load(path) as z:
        state = {k: torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
    dec.load_state_dict(state)
    for p in dec.parameters():
        p.requires_grad_(False)
    return dec.eval()


def load_wfit():
    return load_decoder_npz(WFIT_PATH)


def write_experiment_dir(decoder, directory):
    """Write a DeepSDF experiment dir (specs.json + ModelParameters/latest.pth)
    that `deep_sdf.workspace.config_decoder` (ours or the reference's) loads."""
    import json
    import torch
    os.makedirs(os.path.join(directory, "ModelParameters"), exist_ok=True)
    with open(os.path.join(directory, "specs.json"), "w") as f:
        json.dump({"NetworkArch": "deep_sdf_decoder", "CodeLength": CODE_LEN,
                   "NetworkSpecs": NETWORK_SPECS}, f)
    state = {"module." + k: v.detach().cpu().clone() for k, v in decoder.state_dict().items()}
    torch.save({"model_state_dict": state, "epoch": 0},
               os.path.join(directory, "ModelParameters", "latest.pth"))


# --------------------------------------------------------------------------
# configs (hyper-parameters copied as *values* from the reference's JSON files;
# configs/config_kitti.json:19-39, config_redwood_01053.json, config_freiburg_001.json)
# --------------------------------------------------------------------------
def _cfg(data_type, voxels_dim, k1, k2, k3, k4, b1, b2, iters, s_damp,
         num_depth_samples=50, pose_only_iters=5):
    from reconstruct.utils import ForceKeyErrorDict
    d = dict(
        data_type=data_type, voxels_dim=voxels_dim, DeepSDF_DIR="",
        # detection association (configs/config_kitti.json:14-18; the mono configs share the first three)
        min_bb_area=1600, min_mask_area=1000, downsample_ratio=4.0, num_lidar_max=250, num_lidar_min=10,
        optimizer=dict(
            code_len=CODE_LEN, num_depth_samples=num_depth_samples,
            cut_off_threshold=0.01,
            joint_optim=dict(k1=k1, k2=k2, k3=k3, k4=k4, b1=b1, b2=b2,
                             num_iterations=iters, learning_rate=1.0,
                             scale_damping=s_damp),
        ))
    if data_type == "KITTI":
        d["optimizer"]["pose_only_optim"] = dict(num_iterations=pose_only_iters,
                                                 learning_rate=1.0)
    return ForceKeyErrorDict(**d)


def kitti_config(**kw):
    return _cfg("KITTI", 32, 1.0, 100.0, 0.25, 1e7, 0.20, 0.025, 10, 1.0, **kw)


def redwood_config(iters=20, num_depth_samples=48):
    return _cfg("Redwood", 64, 10.0, 100.0, 2.5, 0.0, 0.20, 0.02, iters, 100.0,
                num_depth_samples=num_depth_samples)


def freiburg_config(iters=20, num_depth_samples=48):
    return _cfg("Freiburg", 64, 1.0, 100.0, 0.5, 0.0, 0.20, 0.025, iters, 100.0,
                num_depth_samples=num_depth_samples)


# --------------------------------------------------------------------------
# detections
# --------------------------------------------------------------------------
def _rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


UPRIGHT = np.diag([1.0, -1.0, -1.0])  # object +y -> camera -y


def _surface_and_rays(rng, T_co, shape, M, n_bg, noise):
    """Camera-facing surface points of the shape (camera frame), their rays, and
    `n_bg` background rays inside the 2-D box but off the silhouette."""
    axes, centres = shape
    R, t = T_co[:3, :3], T_co[:3, 3]
    s = np.cbrt(np.linalg.det(R))
    cam_o = np.linalg.solve(R, -t)          # camera centre in object frame
    pts = []
    need = M
    while need > 0:
        n_c = 6 * need + 32
        prim = rng.integers(0, 2, size=n_c)
        u = rng.normal(size=(n_c, 3))
        u /= np.linalg.norm(u, axis=1, keepdims=True)
        p = u * axes[prim] + centres[prim]              # on primitive `prim`
        f, _ = _prim_f_and_grad(p, axes, centres)
        outside_other = f[np.arange(n_c), 1 - prim] > 0
        n = (p - centres[prim]) / axes[prim] ** 2        # outward normal (unnormalised)
        facing = np.einsum("ij,ij->i", n, cam_o - p) > 0
        p = p[facing & outside_other][:need]
        pts.append(p)
        need -= p.shape[0]
    p_obj = np.concatenate(pts, 0)
    p_cam = p_obj @ R.T + t
    p_cam = p_cam + rng.normal(scale=noise, size=p_cam.shape)
    rays_fg = p_cam / p_cam[:, 2:3]
    depth = p_cam[:, 2].copy()

    # silhouette test: a ray (through the camera centre) hits either primitive?
    def hits(dirs_cam):
        d_obj = np.linalg.solve(R, dirs_cam.T).T
        hit = np.zeros(dirs_cam.shape[0], bool)
        for k in range(2):
            d_o = d_obj / axes[k]
            o_o = (cam_o - centres[k]) / axes[k]
            a = np.einsum("ij,ij->i", d_o, d_o)
            b = 2 * d_o @ o_o
            c = o_o @ o_o - 1.0
            hit |= b * b - 4 * a * c > 0
        return hit

    lo = rays_fg[:, :2].min(0)
    hi = rays_fg[:, :2].max(0)
    pad = 0.08 * (hi - lo)
    lo, hi = lo - pad, hi + pad
    bg = []
    need = n_bg
    while need > 0:
        uv = rng.uniform(lo, hi, size=(4 * need + 16, 2))
        d = np.concatenate([uv, np.ones((uv.shape[0], 1))], 1)
        d = d[~hits(d)][:need]
        bg.append(d)
        need -= d.shape[0]
    rays = np.concatenate([rays_fg] + bg, 0) if n_bg > 0 else rays_fg
    return (p_cam.astype(np.float32), rays.astype(np.float32),
            depth.astype(np.float32), s)


def make_kitti_detection(obj_id, M=250, n_bg=200, noise=0.02, seed0=1234):
    """C1/C2/C5-style detection.  Returns a dict with the
    arguments of ``reconstruct_object`` plus the ground truth."""
    rng = np.random.default_rng(seed0 + obj_id)
    s = rng.uniform(1.8, 2.4)
    yaw = rng.uniform(-np.pi, np.pi)
    t = np.array([rng.uniform(-8, 8), rng.uniform(0.8, 1.4), rng.uniform(8, 30)])
    T_gt = np.eye(4)
    T_gt[:3, :3] = s * _rot_y(yaw) @ UPRIGHT
    T_gt[:3, 3] = t
    code_gt = np.zeros(CODE_LEN)
    code_gt[:3] = 0.3 * rng.normal(size=3)
    shape = shape_params(code_gt)
    pts, rays, depth, _ = _surface_and_rays(rng, T_gt, shape, M, n_bg, noise)
    dyaw = rng.choice([-1.0, 1.0]) * 0.15
    T0 = np.eye(4)
    T0[:3, :3] = 1.1 * s * _rot_y(yaw + dyaw) @ UPRIGHT
    T0[:3, 3] = t + np.array([0.2, 0.05, 0.3])
    return dict(t_cam_obj=T0.astype(np.float32), pts=pts, rays=rays, depth=depth,
                t_cam_obj_gt=T_gt.astype(np.float32), code_gt=code_gt.astype(np.float32))


def make_mono_detection(obj_id, M=256, n_fg=824, n_bg=200, noise=0.01, seed0=4321,
                        scale_range=(0.3, 0.6), dist_range=(1.5, 4.0)):
    """C3-style monocular detection: sparse surface points (SLAM map points),
    `n_fg` foreground rays with depth + `n_bg` background rays, warm-start code."""
    rng = np.random.default_rng(seed0 + obj_id)
    s = rng.uniform(*scale_range)
    yaw = rng.uniform(-np.pi, np.pi)
    dist = rng.uniform(*dist_range)
    t = np.array([rng.uniform(-0.2, 0.2) * dist, rng.uniform(0.0, 0.15) * dist, dist])
    T_gt = np.eye(4)
    T_gt[:3, :3] = s * _rot_y(yaw) @ UPRIGHT
    T_gt[:3, 3] = t
    code_gt = np.zeros(CODE_LEN)
    code_gt[:3] = 0.3 * rng.normal(size=3)
    shape = shape_params(code_gt)
    pts, _, _, _ = _surface_and_rays(rng, T_gt, shape, M, 0, noise * s)
    fg, rays, depth, _ = _surface_and_rays(rng, T_gt, shape, n_fg, n_bg, noise * s)
    dyaw = rng.choice([-1.0, 1.0]) * 0.10
    T0 = np.eye(4)
    T0[:3, :3] = 1.08 * s * _rot_y(yaw + dyaw) @ UPRIGHT
    T0[:3, 3] = t * 1.03
    return dict(t_cam_obj=T0.astype(np.float32), pts=pts, rays=rays, depth=depth,
                code=np.zeros(CODE_LEN, dtype=np.float32),
                t_cam_obj_gt=T_gt.astype(np.float32), code_gt=code_gt.astype(np.float32))


def pose_errors(T_a, T_b):
    """(translation error [m], rotation geodesic [deg], scale ratio) between two Sim3."""
    Ra, Rb = np.asarray(T_a, np.float64)[:3, :3], np.asarray(T_b, np.float64)[:3, :3]
    sa, sb = np.cbrt(np.linalg.det(Ra)), np.cbrt(np.linalg.det(Rb))
    dR = (Ra / sa).T @ (Rb / sb)
    ang = np.degrees(np.arccos(np.clip((np.trace(dR) - 1) / 2, -1, 1)))
    dt = np.linalg.norm(np.asarray(T_a, np.float64)[:3, 3] - np.asarray(T_b, np.float64)[:3, 3])
    return dt, ang, sa / sb

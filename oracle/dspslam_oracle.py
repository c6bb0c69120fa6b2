"""CPU oracle for the DSP-SLAM object-reconstruction hot path.  TEST INFRASTRUCTURE.

A numpy restatement of the reference's algorithm for the path named in
BASELINE.json (`reconstruct/optimizer.py`, `reconstruct/loss.py`,
`reconstruct/loss_utils.py`, `deep_sdf/deep_sdf_decoder.py`).  Every function
cites the reference lines it follows.  Nothing here is product code: only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / reference
arm may import this module; the product path (`dsp-slam-rgbd_b200/`) never
does and fails loudly when the CUDA library is missing.

Parity pinning: the reference has no tests or golden vectors of its own
(SURVEY.md §4), so this oracle is pinned against the *reference's own code
executed in the build container* - `oracle/make_goldens.py` imports
`/root/reference` unmodified (with import stubs) and writes
`tests/golden/*.npz`; `tests/test_oracle_vs_golden.py` checks this file against
those vectors.

`dtype=np.float32` mirrors the reference's arithmetic (everything there is
FP32); `dtype=np.float64` gives a higher-precision truth used to judge which of
two FP32 results is closer.  The reference's autograd Jacobian
(`loss_utils.py:82-113`) is restated as the explicit reverse sweep through the
nine layers, which is what autograd executes.
"""
from __future__ import annotations

import math

import numpy as np

POSE_DIM = 7


# --------------------------------------------------------------------------
# decoder  (deep_sdf/deep_sdf_decoder.py:10-110)
# --------------------------------------------------------------------------
class OracleDecoder:
    """Effective (weight-norm folded) weights of the DeepSDF decoder.

    Weight norm (`deep_sdf_decoder.py:49-54`, torch `weight_norm` default dim=0):
    ``w[o, :] = g[o] * v[o, :] / ||v[o, :]||_2``.
    """

    def __init__(self, state_dict, dtype=np.float32, latent_in=(4,)):
        sd = {k: (v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v))
              for k, v in state_dict.items()}
        self.dtype = dtype
        self.latent_in = tuple(latent_in)
        self.W, self.b = [], []
        k = 0
        while ("lin%d.bias" % k) in sd:
            p = "lin%d." % k
            if p + "weight_g" in sd:
                v = sd[p + "weight_v"].astype(np.float64)
                g = sd[p + "weight_g"].astype(np.float64).reshape(-1, 1)
                w = v * (g / np.linalg.norm(v, axis=1, keepdims=True))
            elif p + "parametrizations.weight.original0" in sd:
                g = sd[p + "parametrizations.weight.original0"].astype(np.float64).reshape(-1, 1)
                v = sd[p + "parametrizations.weight.original1"].astype(np.float64)
                w = v * (g / np.linalg.norm(v, axis=1, keepdims=True))
            else:
                w = sd[p + "weight"].astype(np.float64)
            self.W.append(np.ascontiguousarray(w.astype(dtype)))
            self.b.append(sd[p + "bias"].astype(dtype))
            k += 1
        self.n_layers = k
        self.in_dim = self.W[0].shape[1]
        self.code_len = self.in_dim - 3

    # forward: deep_sdf_decoder.py:75-110 (eval mode: dropout inactive, one final tanh)
    def forward(self, inputs, keep=False):
        x_in = inputs.astype(self.dtype, copy=False)
        h = x_in
        acts = []
        for k in range(self.n_layers):
            if k in self.latent_in:
                h = np.concatenate([h, x_in], axis=-1)
            h = h @ self.W[k].T + self.b[k]
            if k < self.n_layers - 1:
                h = np.maximum(h, 0)
                if keep:
                    acts.append(h > 0)
        out = np.tanh(h)
        return (out, acts) if keep else out

    # decode_sdf: loss_utils.py:51-79
    def sdf(self, code, xyz, max_batch=64 ** 3):
        code = np.asarray(code, self.dtype)
        xyz = np.asarray(xyz, self.dtype)
        outs = []
        for head in range(0, xyz.shape[0], max_batch):
            x = xyz[head:head + max_batch]
            inp = np.concatenate([np.broadcast_to(code, (x.shape[0], code.shape[0])), x], axis=-1)
            outs.append(self.forward(inp)[:, 0])
        return np.concatenate(outs) if outs else np.zeros((0,), self.dtype)

    # get_batch_sdf_jacobian: loss_utils.py:82-113 (out_dim = 1)
    def sdf_jacobian(self, code, xyz, max_batch=1 << 16):
        code = np.asarray(code, self.dtype)
        xyz = np.asarray(xyz, self.dtype)
        n = xyz.shape[0]
        sdf = np.zeros((n,), self.dtype)
        jac = np.zeros((n, self.in_dim), self.dtype)
        for head in range(0, n, max_batch):
            x = xyz[head:head + max_batch]
            inp = np.concatenate([np.broadcast_to(code, (x.shape[0], code.shape[0])), x], axis=-1)
            out, acts = self.forward(inp, keep=True)
            g = (1 - out * out).astype(self.dtype)            # d tanh
            g = g @ self.W[self.n_layers - 1]                 # (n, 512)
            g_in = 0
            for k in range(self.n_layers - 2, -1, -1):
                g = g * acts[k]
                g = g @ self.W[k]
                if k in self.latent_in:                       # cat([hidden, input]) backward
                    split = g.shape[1] - self.in_dim
                    g_in = g_in + g[:, split:]
                    g = g[:, :split]
            sdf[head:head + x.shape[0]] = out[:, 0]
            jac[head:head + x.shape[0]] = g + g_in
        return sdf, jac


# --------------------------------------------------------------------------
# Lie-group helpers  (loss_utils.py:117-243)
# --------------------------------------------------------------------------
def _hat(w, dtype):
    return np.array([[0., -w[2], w[1]], [w[2], 0., -w[0]], [-w[1], w[0], 0.]], dtype=dtype)


def exp_se3(x, dtype=np.float32):
    """loss_utils.py:139-173; ordering (v, w)."""
    x = np.asarray(x, dtype)
    v, w = x[:3], x[3:6]
    w_hat = _hat(w, dtype)
    w_hat2 = w_hat @ w_hat
    theta = dtype(np.sqrt(np.sum(w * w, dtype=dtype)))
    eye = np.eye(3, dtype=dtype)
    if theta <= 1e-8:
        e_w, j = eye, eye
    else:
        st, ct = dtype(np.sin(theta)), dtype(np.cos(theta))
        th2, th3 = theta ** 2, theta ** 3
        e_w = eye + w_hat * st / theta + w_hat2 * (dtype(1.) - ct) / th2
        k1 = (1 - ct) / th2
        k2 = (theta - st) / th3
        j = eye + k1 * w_hat + k2 * w_hat2
    T = np.eye(4, dtype=dtype)
    T[:3, :3] = e_w
    T[:3, 3] = j.astype(dtype) @ v
    return T


def exp_sim3(x, dtype=np.float32):
    """loss_utils.py:198-243; ordering (v, w, sigma).  Reproduces the quirk
    ``c = 0 if s <= eps`` for negative sigma when theta > 1e-8 (`:233`)."""
    x = np.asarray(x, dtype)
    v, w, s = x[:3], x[3:6], x[6]
    w_hat = _hat(w, dtype)
    w_hat2 = w_hat @ w_hat
    theta = dtype(np.sqrt(np.sum(w * w, dtype=dtype)))
    th2 = theta ** 2
    st, ct = dtype(np.sin(theta)), dtype(np.cos(theta))
    e_s = dtype(np.exp(s))
    s2 = s ** 2
    eye = np.eye(3, dtype=dtype)
    eps = 1e-8
    if theta <= 1e-8:
        e_w = eye
        j = eye if s == 0 else ((e_s - dtype(1.)) / s) * eye
    else:
        e_w = eye + w_hat * st / theta + w_hat2 * (dtype(1.) - ct) / th2
        a = e_s * st
        b = e_s * ct
        c = dtype(0.) if s <= eps else (e_s - dtype(1.)) / s
        k1 = (a * s + (1 - b) * theta) / (s2 + th2)
        k2 = c - ((b - 1) * s + a * theta) / (s2 + th2)
        j = c * eye + k1 * w_hat / theta + k2 * w_hat2 / th2
    T = np.eye(4, dtype=dtype)
    T[:3, :3] = e_s * e_w
    T[:3, 3] = np.asarray(j, dtype) @ v
    return T


def points_to_pose_jacobian_sim3(p):
    """loss_utils.py:176-195: [I | -hat(x) | x]  -> (n, 3, 7)."""
    n = p.shape[0]
    J = np.zeros((n, 3, 7), p.dtype)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    J[:, 0, 0] = J[:, 1, 1] = J[:, 2, 2] = 1
    J[:, 0, 4], J[:, 0, 5] = z, -y
    J[:, 1, 3], J[:, 1, 5] = -z, x
    J[:, 2, 3], J[:, 2, 4] = y, -x
    J[:, :, 6] = p
    return J


# --------------------------------------------------------------------------
# robust kernel  (loss_utils.py:246-275)
# --------------------------------------------------------------------------
def get_robust_res(res, b):
    res = res.reshape(-1)
    x = np.abs(res)
    rn = np.where(x <= b, x * x, 2 * res.dtype.type(b) * x - res.dtype.type(b) ** 2).astype(res.dtype)
    x = np.where(x == 0, res.dtype.type(1), x)
    w = np.sqrt(rn) / x
    robust = w * res
    with np.errstate(invalid="ignore", divide="ignore"):
        loss = np.mean(robust ** 2) if robust.size else res.dtype.type(np.nan)
    return robust, loss, w


# --------------------------------------------------------------------------
# residual terms  (loss.py)
# --------------------------------------------------------------------------
def transform_points(T, p):
    """loss.py:31-32 / :76-77:  (p[..., None, :] * R).sum(-1) + t  ==  p @ R^T + t."""
    return (p[..., None, :] * T[:3, :3]).sum(-1) + T[:3, 3]


def compute_sdf_loss(dec, pts_cam, T_oc, code):
    """loss.py:22-43 -> jac_pose (M,7), jac_code (M,L), res (M,)."""
    dt = dec.dtype
    p_obj = transform_points(T_oc.astype(dt), pts_cam.astype(dt))
    sdf, de_di = dec.sdf_jacobian(code, p_obj)
    de_dx = de_di[:, -3:]
    J = points_to_pose_jacobian_sim3(p_obj)
    jac_pose = np.einsum("ni,nij->nj", de_dx, J)
    return jac_pose, de_di[:, :-3], sdf


def linspace_torch(a, b, n, dtype=np.float32):
    """torch.linspace's symmetric evaluation (start + i*step for the lower half,
    end - (n-1-i)*step for the upper half), evaluated in double and rounded once,
    which is what torch's CPU kernel returns for a float32 result."""
    a, b = float(dtype(a)), float(dtype(b))
    step = (b - a) / (n - 1)
    i = np.arange(n)
    return np.where(i < n // 2, a + step * i, b - step * (n - 1 - i)).astype(dtype)


def compute_render_loss(dec, rays, depth_obs, T_oc, sampled_depth, code, th=0.01,
                        return_aux=False):
    """loss.py:60-166.  Returns None when fewer than 10 samples fall inside the
    unit ball (`:87-88`); otherwise (jac_pose (K,7), jac_code (K,L), res (K,)).
    Rows are ordered by (ray, sample) like the reference's boolean-mask gathers."""
    dt = dec.dtype
    rays = rays.astype(dt)
    sd = sampled_depth.astype(dt)
    T = T_oc.astype(dt)
    N, D = rays.shape[0], sd.shape[0]
    pts_cam = rays[:, None, :] * sd[:, None]                      # :74
    pts_obj = transform_points(T, pts_cam)                        # :76-77
    valid = np.sqrt((pts_obj * pts_obj).sum(-1)) < 1.0            # :82
    n_valid = int(valid.sum())
    if n_valid < 10:
        return None
    sdf = dec.sdf(code, pts_obj[valid])                           # :91-92
    th_t = dt(th)
    occ = np.zeros((N, D), dt)
    occ[valid] = dt(0.5) - np.clip(sdf, -th_t, th_t) / (2 * th_t)  # :98-100, loss_utils:40-48
    wg = np.zeros((N, D), bool)
    wg[valid] = (sdf > -th_t) & (sdf < th_t)                      # :102
    d_min, d_max = sd[0], sd[-1]
    acc = np.cumprod(1 - occ, axis=-1, dtype=dt)                  # :113
    acc_aug = np.concatenate([np.ones((N, 1), dt), acc], -1)
    o = np.concatenate([occ, np.ones((N, 1), dt)], -1)
    d = np.concatenate([sd, np.array([dt(1.1) * d_max], dt)])
    term = o * acc_aug                                            # :126
    d_u = (d * term).sum(-1, dtype=dt)                            # :128
    # de_do_k = sum_{j>=k} acc_j / (1 - o_k)                       :132-136
    suffix = np.cumsum(acc[:, ::-1], axis=-1, dtype=dt)[:, ::-1]
    with np.errstate(divide="ignore", invalid="ignore"):
        de_do = suffix / (1 - occ)
    sel = wg & (de_do > 1e-2)                                     # :139
    ri, si = np.nonzero(sel)
    delta_d = (d_max - d_min) / dt(D - 1)                         # :142
    do_ds = dt(-1. / (2 * th))
    de_ds = de_do[ri, si] * delta_d * do_ds                       # :144
    res = depth_obs.astype(dt)[ri] - d_u[ri]                      # :149-150
    res = np.clip(res, dt(-0.30), dt(0.30))                       # :153-154
    p = pts_obj[ri, si]
    _, ds_di = dec.sdf_jacobian(code, p)                          # :157-158
    de_di = de_ds[:, None] * ds_di
    J = points_to_pose_jacobian_sim3(p)
    jac_pose = np.einsum("ni,nij->nj", de_di[:, -3:], J)          # :160-163
    out = (jac_pose, de_di[:, :-3], res)
    if return_aux:
        return out + (dict(n_valid=n_valid, rows=(ri, si), d_u=d_u, occ=occ, de_do=de_do,
                           pts=p, de_ds=de_ds, m_with_grad=int(wg.sum())),)
    return out


def compute_rotation_loss_sim3(T_oc, dtype=np.float32):
    """loss.py:169-192 (CPU float32 in the reference)."""
    T_co = np.linalg.inv(T_oc.astype(dtype))
    r_co = T_co[:3, :3]
    scale = dtype(np.linalg.det(r_co)) ** dtype(1 / 3)
    r_co = r_co / scale
    r_oc = np.linalg.inv(r_co)
    ey = np.array([0., 1., 0.], dtype)
    ng = np.array([0., -1., 0.], dtype)
    res = dtype(1.) - (r_co @ ey) @ ng
    if res < 1e-7:
        return np.zeros(7, dtype), dtype(0.)
    J = np.zeros(7, dtype)
    J[3:6] = np.cross(r_oc @ ng, ey)
    return J, dtype(res)


# --------------------------------------------------------------------------
# Gauss-Newton drivers  (optimizer.py)
# --------------------------------------------------------------------------
class OracleOptimizer:
    """optimizer.py:26-213 restated on numpy."""

    def __init__(self, dec, configs):
        self.dec = dec
        oc = configs.optimizer
        j = oc.joint_optim
        self.k1, self.k2, self.k3, self.k4 = j.k1, j.k2, j.k3, j.k4
        self.b1, self.b2 = j.b1, j.b2
        self.lr = j.learning_rate
        self.s_damp = j.scale_damping
        self.num_iterations_joint_optim = j.num_iterations
        self.code_len = oc.code_len
        self.num_depth_samples = oc.num_depth_samples
        self.cut_off = oc.cut_off_threshold
        if configs.data_type == "KITTI":
            self.num_iterations_pose_only = oc.pose_only_optim.num_iterations

    # optimizer.py:46-87
    def estimate_pose_cam_obj(self, t_co_se3, scale, pts, code, trace=None):
        dt = self.dec.dtype
        T_co = np.array(t_co_se3, dtype=dt)
        T_co[:3, :3] *= dt(scale)
        T_oc = np.linalg.inv(T_co)
        code = np.asarray(code, dt)
        pts = np.asarray(pts, dt)
        for e in range(self.num_iterations_pose_only):
            jp, _, res = compute_sdf_loss(self.dec, pts, T_oc, code)
            j = jp[:, :6]
            n = dt(j.shape[0])
            H = (j.T @ j) / n + dt(1e-2) * np.eye(6, dtype=dt)
            b = -(j.T @ res) / n
            dx = np.linalg.inv(H) @ b
            if trace is not None:
                trace.append(dict(H=H.copy(), b=b.copy(), T_oc=T_oc.copy()))
            T_oc = exp_se3(dx, dt) @ T_oc
            if e == 4:
                pts = pts[np.abs(res) <= 0.05]
        T_co = np.linalg.inv(T_oc)
        T_co[:3, :3] /= dt(scale)
        return T_co

    # optimizer.py:90-205
    def reconstruct_object(self, t_cam_obj, pts, rays, depth, code=None, trace=None):
        dt = self.dec.dtype
        L = self.code_len
        z = np.zeros(L, dt) if code is None else np.array(code[:L], dtype=dt)
        T_oc = np.linalg.inv(np.asarray(t_cam_obj, dt))
        rays = np.asarray(rays, dt)
        n_fg = depth.shape[0]
        depth_obs = np.concatenate([np.asarray(depth, dt), np.zeros(rays.shape[0] - n_fg, dt)])
        pts = np.asarray(pts, dt)
        loss = 0.
        fail = dict(t_cam_obj=None, code=None, is_good=False)
        for e in range(self.num_iterations_joint_optim):
            T_co = np.linalg.inv(T_oc)
            scale = dt(np.linalg.det(T_co[:3, :3])) ** dt(1 / 3)
            dmin, dmax = T_co[2, 3] - dt(1.0) * scale, T_co[2, 3] + dt(1.0) * scale
            sd = linspace_torch(dmin, dmax, self.num_depth_samples, dt)
            depth_obs[n_fg:] = dt(1.1) * dmax
            jp_s, jc_s, res_s = compute_sdf_loss(self.dec, pts, T_oc, z)
            rr_s, sdf_loss, _ = get_robust_res(res_s, self.b2)
            if math.isnan(sdf_loss):
                return dict(fail, loss=loss)
            rnd = compute_render_loss(self.dec, rays, depth_obs, T_oc, sd, z, th=self.cut_off,
                                      return_aux=True)
            if rnd is None:
                return dict(fail, loss=loss)
            jp_r, jc_r, res_r, aux = rnd
            rr_r, render_loss, _ = get_robust_res(res_r, self.b1)
            if math.isnan(render_loss):
                return dict(fail, loss=loss)
            J_rot, res_rot = compute_rotation_loss_sim3(T_oc, dt)
            loss = self.k1 * render_loss + self.k2 * sdf_loss

            Js = np.concatenate([jp_s, jc_s], -1)
            H_s = dt(self.k2) * (Js.T @ Js) / dt(Js.shape[0])
            b_s = -dt(self.k2) * (Js.T @ rr_s) / dt(Js.shape[0])
            Jr = np.concatenate([jp_r, jc_r], -1)
            H_r = dt(self.k1) * (Jr.T @ Jr) / dt(Jr.shape[0])
            b_r = -dt(self.k1) * (Jr.T @ rr_r) / dt(Jr.shape[0])
            H = H_r + H_s
            H[7:7 + L, 7:7 + L] += dt(self.k3) * np.eye(L, dtype=dt)
            b = b_r + b_s
            b[7:7 + L] -= dt(self.k3) * z
            H[:7, :7] += dt(self.k4) * np.outer(J_rot, J_rot)
            b_rot = -(J_rot * res_rot)
            b[:7] -= dt(self.k4) * b_rot
            H[:7, :7] += np.eye(7, dtype=dt)
            H[6, 6] += dt(self.s_damp)
            dx = np.linalg.inv(H) @ b
            if trace is not None:
                trace.append(dict(H=H.copy(), b=b.copy(), T_oc=T_oc.copy(), code=z.copy(),
                                  K=Jr.shape[0], n_valid=aux["n_valid"],
                                  sdf_loss=sdf_loss, render_loss=render_loss, loss=loss))
            T_oc = exp_sim3(dt(self.lr) * dx[:7], dt) @ T_oc
            z = z + dt(self.lr) * dx[7:7 + L]
        return dict(t_cam_obj=np.linalg.inv(T_oc), code=z, is_good=True, loss=loss)

    # optimizer.py:207-213 + loss.py:46-56
    def compute_sdf_loss_objectpoint_zhjd(self, pts_obj, code):
        sdf, _ = self.dec.sdf_jacobian(np.asarray(code[:self.code_len], self.dec.dtype),
                                       np.asarray(pts_obj, self.dec.dtype))
        return sdf.mean(dtype=self.dec.dtype)


# --------------------------------------------------------------------------
# mesh grid  (optimizer.py:216-233, utils.py:97-116)
# --------------------------------------------------------------------------
def create_voxel_grid(vol_dim, coord_mode="reference_shear"):
    """utils.py:97-116.  Under torch>=1.6 ``index / vol_dim`` is a true division,
    so the x/y 'indices' keep a fractional part (sheared grid)."""
    n = vol_dim ** 3
    i = np.arange(n, dtype=np.int64)
    vs = np.float32(2.0 / (vol_dim - 1))
    G = np.float32(vol_dim)
    out = np.empty((n, 3), np.float32)
    z = (i % vol_dim).astype(np.float32)
    if coord_mode == "reference_shear":
        q = i.astype(np.float32) / G
        y = np.fmod(q, G)
        x = np.fmod(q / G, G)
    else:
        y = ((i // vol_dim) % vol_dim).astype(np.float32)
        x = ((i // vol_dim ** 2) % vol_dim).astype(np.float32)
    out[:, 0] = x * vs + np.float32(-1)
    out[:, 1] = y * vs + np.float32(-1)
    out[:, 2] = z * vs + np.float32(-1)
    return out


def grid_sdf(dec, code, vol_dim, coord_mode="reference_shear"):
    """optimizer.py:224-227: decode_sdf over the voxel grid -> (G, G, G), x slowest."""
    pts = create_voxel_grid(vol_dim, coord_mode)
    return dec.sdf(np.asarray(code[:dec.code_len]), pts).reshape(vol_dim, vol_dim, vol_dim)

"""Stock-PyTorch FP32 restatement of `Optimizer.reconstruct_object` that runs on the GPU.

A like-for-like GPU data point for SURVEY.md section 8(d) ("the reference on the B200 itself"): the reference's own
files cannot travel to the GPU box, so this is a PORT of its loop (optimizer.py:90-205, loss.py:22-166,
loss_utils.py:40-275) issued the way the reference issues it - the weight-normed `nn.Module` decoder, one shared code
expanded and concatenated per query row, `no_grad` forward for the ray samples, forward + one reverse-mode pass for
the Jacobians, boolean-mask gathers, `torch.inverse` of the 71x71 system, one detection at a time - with every tensor
on `cuda` and TF32 off.  Only `tools/bench_configs.py torchgn` uses it; it is neither product code nor the oracle.
"""
import torch


def _hat(w):
    z = torch.zeros((), dtype=w.dtype, device=w.device)
    return torch.stack([torch.stack([z, -w[2], w[1]]), torch.stack([w[2], z, -w[0]]), torch.stack([-w[1], w[0], z])])


def exp_sim3(x):
    """loss_utils.py:198-243, ordering (v, w, sigma), including `c = 0 if sigma <= eps`."""
    v, w, s = x[:3], x[3:6], x[6]
    wh = _hat(w)
    wh2 = wh @ wh
    th = torch.sqrt((w * w).sum())
    eye = torch.eye(3, dtype=x.dtype, device=x.device)
    es = torch.exp(s)
    if float(th) <= 1e-8:
        ew = eye
        j = eye if float(s) == 0 else ((es - 1) / s) * eye
    else:
        st, ct = torch.sin(th), torch.cos(th)
        ew = eye + wh * st / th + wh2 * (1 - ct) / th ** 2
        a, b = es * st, es * ct
        c = torch.zeros_like(es) if float(s) <= 1e-8 else (es - 1) / s
        k1 = (a * s + (1 - b) * th) / (s ** 2 + th ** 2)
        k2 = c - ((b - 1) * s + a * th) / (s ** 2 + th ** 2)
        j = c * eye + k1 * wh / th + k2 * wh2 / th ** 2
    T = torch.eye(4, dtype=x.dtype, device=x.device)
    T[:3, :3] = es * ew
    T[:3, 3] = j @ v
    return T


def pose_jacobian(p):
    """loss_utils.py:176-195: [I | -hat(p) | p] per point."""
    J = torch.zeros((p.shape[0], 3, 7), dtype=p.dtype, device=p.device)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    J[:, 0, 0] = J[:, 1, 1] = J[:, 2, 2] = 1
    J[:, 0, 4], J[:, 0, 5] = z, -y
    J[:, 1, 3], J[:, 1, 5] = -z, x
    J[:, 2, 3], J[:, 2, 4] = y, -x
    J[:, :, 6] = p
    return J


def robust(res, b):
    """loss_utils.py:246-275."""
    x = res.abs()
    rn = torch.where(x <= b, x * x, 2 * b * x - b * b)
    w = rn.sqrt() / torch.where(x == 0, torch.ones_like(x), x)
    r = w * res
    return r, (r * r).mean()


def decode_sdf(net, code, xyz, max_batch=64 ** 3):
    outs = []
    with torch.no_grad():
        for head in range(0, xyz.shape[0], max_batch):
            x = xyz[head: head + max_batch]
            outs.append(net(torch.cat([code.expand(x.shape[0], -1), x], dim=1)).squeeze(1))
    return torch.cat(outs)


def sdf_jacobian(net, code, xyz):
    inp = torch.cat([code.expand(xyz.shape[0], -1), xyz], dim=1).unsqueeze(1).requires_grad_(True)
    y = net(inp)
    y.backward(torch.ones_like(y))
    return y.detach().reshape(-1), inp.grad.reshape(xyz.shape[0], -1)


def xform(T, p):
    return (p[..., None, :] * T[:3, :3]).sum(-1) + T[:3, 3]


def reconstruct_object(net, cfg, t_cam_obj, pts, rays, depth):
    """One detection, all tensors on the decoder's device.  Returns (T_cam_obj 4x4, code, is_good, loss)."""
    dev = next(net.parameters()).device
    j = cfg.optimizer.joint_optim
    L, D, th = cfg.optimizer.code_len, cfg.optimizer.num_depth_samples, cfg.optimizer.cut_off_threshold
    f = dict(dtype=torch.float32, device=dev)
    z = torch.zeros(L, **f)
    T_oc = torch.inverse(torch.as_tensor(t_cam_obj, **f))
    pts, rays, depth = (torch.as_tensor(a, **f) for a in (pts, rays, depth))
    n_fg = depth.shape[0]
    depth_obs = torch.cat([depth, torch.zeros(rays.shape[0] - n_fg, **f)])
    loss = 0.0
    for _ in range(j.num_iterations):
        T_co = torch.inverse(T_oc)
        scale = torch.det(T_co[:3, :3]) ** (1 / 3)
        dmin, dmax = T_co[2, 3] - scale, T_co[2, 3] + scale
        sd = torch.linspace(float(dmin), float(dmax), D, **f)
        depth_obs[n_fg:] = 1.1 * dmax
        # SDF term (loss.py:22-43)
        p_obj = xform(T_oc, pts)
        res_s, di = sdf_jacobian(net, z, p_obj)
        jp_s = torch.einsum("ni,nij->nj", di[:, -3:], pose_jacobian(p_obj))
        Js = torch.cat([jp_s, di[:, :-3]], -1)
        rr_s, sdf_loss = robust(res_s, j.b2)
        # render term (loss.py:60-166)
        pc = rays[:, None, :] * sd[:, None]
        po = xform(T_oc, pc)
        valid = po.norm(dim=-1) < 1.0
        if int(valid.sum()) < 10:
            return None, None, False, loss
        sdf = decode_sdf(net, z, po[valid])
        occ = torch.zeros(valid.shape, **f)
        occ[valid] = 0.5 - sdf.clamp(-th, th) / (2 * th)
        wg = torch.zeros_like(valid)
        wg[valid] = (sdf > -th) & (sdf < th)
        acc = torch.cumprod(1 - occ, dim=-1)
        ones = torch.ones((rays.shape[0], 1), **f)
        term = torch.cat([occ, ones], -1) * torch.cat([ones, acc], -1)
        d_u = (torch.cat([sd, 1.1 * sd[-1:]]) * term).sum(-1)
        de_do = acc.flip(-1).cumsum(-1).flip(-1) / (1 - occ)
        sel = wg & (de_do > 1e-2)
        ri, si = torch.nonzero(sel, as_tuple=True)
        de_ds = de_do[ri, si] * ((sd[-1] - sd[0]) / (D - 1)) * (-1.0 / (2 * th))
        res_r = (depth_obs[ri] - d_u[ri]).clamp(-0.30, 0.30)
        p = po[ri, si]
        _, ds_di = sdf_jacobian(net, z, p)
        de_di = de_ds[:, None] * ds_di
        Jr = torch.cat([torch.einsum("ni,nij->nj", de_di[:, -3:], pose_jacobian(p)), de_di[:, :-3]], -1)
        rr_r, render_loss = robust(res_r, j.b1)
        if bool(torch.isnan(sdf_loss)) or bool(torch.isnan(render_loss)):
            return None, None, False, loss
        loss = float(j.k1 * render_loss + j.k2 * sdf_loss)
        # rotation prior (loss.py:169-192)
        r_co = T_co[:3, :3] / scale
        ey, ng = torch.tensor([0., 1., 0.], **f), torch.tensor([0., -1., 0.], **f)
        res_rot = 1.0 - (r_co @ ey) @ ng
        J_rot = torch.zeros(7, **f)
        if float(res_rot) >= 1e-7:
            J_rot[3:6] = torch.linalg.cross(torch.inverse(r_co) @ ng, ey)
        else:
            res_rot = torch.zeros((), **f)
        # normal equations (optimizer.py:150-186)
        H = j.k1 * (Jr.T @ Jr) / Jr.shape[0] + j.k2 * (Js.T @ Js) / Js.shape[0]
        b = -j.k1 * (Jr.T @ rr_r) / Jr.shape[0] - j.k2 * (Js.T @ rr_s) / Js.shape[0]
        H[7:, 7:] += j.k3 * torch.eye(L, **f)
        b[7:] -= j.k3 * z
        H[:7, :7] += j.k4 * torch.outer(J_rot, J_rot) + torch.eye(7, **f)
        b[:7] += j.k4 * J_rot * res_rot
        H[6, 6] += j.scale_damping
        dx = torch.inverse(H) @ b
        T_oc = exp_sim3(j.learning_rate * dx[:7]) @ T_oc
        z = z + j.learning_rate * dx[7:]
    return torch.inverse(T_oc), z, True, loss

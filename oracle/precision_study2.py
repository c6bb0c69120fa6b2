"""TEST INFRASTRUCTURE (CPU, numpy): which operand roundings of the tensor-core kernels keep a teacher-forced
Gauss-Newton step inside north_star's per-step tolerances (1 cm / 0.1 deg / 1e-2 relative code)?

For every reference golden object and iteration e the oracle performs ONE step from the reference's own state
(T_oc[e], code[e]) with an emulation of the FP16-operand kernels plugged in separately for
  F: the forward-only evaluation of the ray samples      (k_tc<3>, `decode_sdf`, loss.py:91-92)
  J: the SDF + Jacobian evaluation of the selected rows   (k_tc<5>, `get_batch_sdf_jacobian`, loss_utils.py:82-113)
and the result is compared with the reference's state at e + 1.  Roundings: w weights, a forward activations,
g backward gradients to FP16; W / A / G the same operands as hi + lo FP16 pairs (two MMAs per K step each).

Usage: python oracle/precision_study2.py [n_objects] [variant ...]"""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
from dspslam import synth  # noqa: E402
from oracle import dspslam_oracle as orc  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")


def r16(x):
    return x.astype(np.float16).astype(np.float32)


def r16x2(x):
    h = r16(x)
    return h + r16(x - h)


RND = {"": lambda x: x, "1": r16, "2": r16x2}


class EmuDecoder(orc.OracleDecoder):
    """spec = (fw, fa, jw, ja, jg), each '' (FP32), '1' (FP16) or '2' (hi + lo FP16)."""

    def __init__(self, sd, spec):
        super().__init__(sd)
        self.spec = spec
        need = set((spec[0], spec[2]) + ((spec[7],) if len(spec) > 5 else ()))
        if need & {"f", "b", "lo3", "hi4"}:
            need = (need - {"f", "b", "lo3", "hi4"}) | {"1", "2"}
        self.Wv = {m: [RND[m](w) for w in self.W] for m in need}

    def _fwd(self, inp, wm, am, keep):
        if wm in ("lo3", "hi4"):      # hi+lo weights in layers 1-3 / 4-7 only
            W = [self.Wv["2"][k] if ((k <= 3) == (wm == "lo3")) else self.Wv["1"][k] for k in range(self.n_layers)]
        else:
            W = self.Wv[wm]
        h = inp
        acts = []
        nl = self.n_layers
        for k in range(nl):
            if k == 0 or k == nl - 1:
                h = h @ self.W[k].T + self.b[k]              # layer 0: hi/lo extras MMA; layer 8: FP32 epilogue
            elif k in self.latent_in:
                split = W[k].shape[1] - self.in_dim
                h = h @ W[k][:, :split].T + inp @ self.W[k][:, split:].T + self.b[k]
            else:
                h = h @ W[k].T + self.b[k]
            if k < nl - 1:
                if keep:
                    acts.append(h > 0)
                h = np.maximum(h, 0)
                if k < nl - 2:
                    h = RND[am](h)
        return np.tanh(h), acts

    def sdf(self, code, xyz, max_batch=64 ** 3):
        code = np.asarray(code, np.float32)
        x = np.asarray(xyz, np.float32)
        if x.shape[0] == 0:
            return np.zeros((0,), np.float32)
        inp = np.concatenate([np.broadcast_to(code, (x.shape[0], code.shape[0])), x], axis=-1)
        s = self._fwd(inp, self.spec[0], self.spec[1], False)[0][:, 0]
        if len(self.spec) > 5:
            # second, FP32-grade evaluation of the rows the FP16 pass leaves within eps of a decision threshold
            # ('edge': | |s| - th | < eps) or anywhere inside the widened band ('band': |s| < th + eps)
            kind, eps, th = self.spec[5], self.spec[6], 0.01
            m = (np.abs(np.abs(s) - th) < eps) if kind == "edge" else (np.abs(s) < th + eps)
            self.refined = getattr(self, "refined", 0) + int(m.sum())
            self.total = getattr(self, "total", 0) + int(m.size)
            if m.any():
                s = s.copy()
                s[m] = self._fwd(inp[m], self.spec[7], self.spec[8], False)[0][:, 0]
        return s

    def sdf_jacobian(self, code, xyz, max_batch=1 << 16):
        code = np.asarray(code, np.float32)
        x = np.asarray(xyz, np.float32)
        inp = np.concatenate([np.broadcast_to(code, (x.shape[0], code.shape[0])), x], axis=-1)
        jw = self.spec[2]
        out, acts = self._fwd(inp, "1" if jw == "b" else ("2" if jw == "f" else jw), self.spec[3], True)
        W = self.Wv["2" if jw == "b" else ("1" if jw in ("f", "lo3", "hi4") else jw)]      # 'f' / 'b': hi+lo weights in the forward / backward sweep only
        gm = RND[self.spec[4]]
        g = (1 - out * out).astype(np.float32) @ self.W[self.n_layers - 1]
        g_in = 0
        for k in range(self.n_layers - 2, -1, -1):
            g = gm(g * acts[k])
            g = g @ W[k]
            if k in self.latent_in:
                split = g.shape[1] - self.in_dim
                g_in = g_in + g[:, split:]
                g = g[:, :split]
        return out[:, 0], (g + g_in).astype(np.float32)


VARIANTS = {
    "fp32": ("", "", "", "", ""),
    "tc (w a | w a g)": ("1", "1", "1", "1", "1"),
    "F tc | J fp32": ("1", "1", "", "", ""),
    "F fp32 | J tc": ("", "", "1", "1", "1"),
    "F tc | J W a g": ("1", "1", "2", "1", "1"),
    "F tc | J W A G": ("1", "1", "2", "2", "2"),
    "F W a | J W a g": ("2", "1", "2", "1", "1"),
    "F W A | J W A G": ("2", "2", "2", "2", "2"),
    "F W a | J W A G": ("2", "1", "2", "2", "2"),
    "F fp32 | J W a g": ("", "", "2", "1", "1"),
    "F fp32 | J Wfwd a g": ("", "", "f", "1", "1"),
    "F fp32 | J Wfwd(1-3) a g": ("", "", "lo3", "1", "1"),
    "F fp32 | J Wfwd(4-7) a g": ("", "", "hi4", "1", "1"),
    "F fp32 | J Wbwd a g": ("", "", "b", "1", "1"),
    "F fp32 | J w A G": ("", "", "1", "2", "2"),
    "F fp32 | J W A g": ("", "", "2", "2", "1"),
    "F tc+edge5e-4 | J fp32": ("1", "1", "", "", "", "edge", 5e-4, "", ""),
    "F tc+band5e-4 | J fp32": ("1", "1", "", "", "", "band", 5e-4, "", ""),
    "F tc+edge5e-4(WA) | J tc": ("1", "1", "1", "1", "1", "edge", 5e-4, "2", "2"),
    "F tc+edge5e-4(WA) | J W a g": ("1", "1", "2", "1", "1", "edge", 5e-4, "2", "2"),
    "F tc+band5e-4(WA) | J W a g": ("1", "1", "2", "1", "1", "band", 5e-4, "2", "2"),
    "F tc+band5e-4(WA) | J tc": ("1", "1", "1", "1", "1", "band", 5e-4, "2", "2"),
    "F tc+band5e-4(WA) | J W A G": ("1", "1", "2", "2", "2", "band", 5e-4, "2", "2"),
}


def objects(n):
    """(name, golden dict, detection, config, n_iterations)"""
    out = []
    bulk = os.path.join(G, "bulk_c1.npz")
    if os.path.exists(bulk):
        with np.load(bulk) as z:
            b = {k: z[k] for k in z.files}
        for k in range(min(n, b["T_oc"].shape[0])):
            if not b["is_good"][k]:
                continue
            g = dict(T_oc=b["T_oc"][k], code_it=b["code_it"][k], t_cam_obj=b["t_cam_obj"][k], code=b["code"][k])
            out.append(("bulk_c1[%d]" % k, g, synth.make_kitti_detection(k), synth.kitti_config()))
    else:
        for k in range(min(n, 4)):
            with np.load(os.path.join(G, "c1_obj%d.npz" % k)) as z:
                g = {kk: z[kk] for kk in z.files}
            det = dict(pts=g["in_pts"], rays=g["in_rays"], depth=g["in_depth"])
            out.append(("c1_obj%d" % k, g, det, synth.kitti_config()))
    return out


def step_errors(dec, cfg, g, det, e):
    cfg.optimizer.joint_optim.num_iterations = 1
    T_co = np.linalg.inv(g["T_oc"][e].astype(np.float64)).astype(np.float32)
    o = orc.OracleOptimizer(dec, cfg).reconstruct_object(T_co, det["pts"], det["rays"], det["depth"], code=g["code_it"][e])
    n_it = g["T_oc"].shape[0]
    Tn, cn = ((np.linalg.inv(g["T_oc"][e + 1].astype(np.float64)), g["code_it"][e + 1]) if e + 1 < n_it
              else (g["t_cam_obj"], g["code"]))
    if not o["is_good"]:
        return 1., 90., 1.
    dt, ang, _ = synth.pose_errors(o["t_cam_obj"], Tn)
    ec = np.linalg.norm(o["code"] - cn) / max(np.linalg.norm(cn), 0.1)
    return dt, ang, ec


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    names = sys.argv[2:] or list(VARIANTS)
    sd = synth.load_wfit().state_dict()
    objs = objects(n)
    print("%d objects" % len(objs), flush=True)
    for name in names:
        dec = EmuDecoder(sd, VARIANTS[name])
        errs = []
        for _, g, det, cfg in objs:
            for e in range(g["T_oc"].shape[0]):
                errs.append(step_errors(dec, cfg, g, det, e))
        E = np.array(errs).reshape(len(objs), -1, 3)
        ok = (E[..., 0] < 1e-2) & (E[..., 1] < 0.1) & (E[..., 2] < 1e-2)
        print("%-22s steps ok %.3f | objects with every step ok %.3f | dt max %.1e  rot med %.1e q95 %.1e max %.1e  "
              "code med %.1e q95 %.1e max %.1e" % (name, ok.mean(), ok.all(axis=1).mean(), E[..., 0].max(),
                                                   np.median(E[..., 1]), np.quantile(E[..., 1], .95), E[..., 1].max(),
                                                   np.median(E[..., 2]), np.quantile(E[..., 2], .95), E[..., 2].max()),
              flush=True)
        if hasattr(dec, "refined"):
            print("    refined rows: %d of %d forward rows (%.2f %%)" % (dec.refined, dec.total, 100. * dec.refined / dec.total), flush=True)


if __name__ == "__main__":
    main()

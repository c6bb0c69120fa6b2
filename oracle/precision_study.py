"""TEST INFRASTRUCTURE (CPU, numpy): which roundings of the tensor-core path move a Gauss-Newton step?

Emulates, inside the oracle's SDF+Jacobian evaluation only (the ray samples stay FP32), FP16 rounding of
  w: the weights,  a: the forward activations,  g: the backward gradients
in any combination and prints the relative change of the first GN step's code / pose update against the
all-FP32 oracle over a few synthetic C1 detections.   Usage: python oracle/precision_study.py [n_objects]"""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
from dspslam import synth  # noqa: E402
from oracle import dspslam_oracle as orc  # noqa: E402


def r16(x):
    return x.astype(np.float16).astype(np.float32)


class QuantDecoder(orc.OracleDecoder):
    def __init__(self, sd, qw=False, qa=False, qg=False, split_a=False, split_w=False):
        super().__init__(sd)
        self.qw, self.qa, self.qg, self.split_a, self.split_w = qw, qa, qg, split_a, split_w
        self.Wq = [r16(w) if qw else w for w in self.W]
        if split_w:                      # hi + lo FP16 weights: ~22 bits
            self.Wq = [r16(w) + r16(w - r16(w)) for w in self.W]

    def _act(self, h):
        if self.split_a:
            return r16(h) + r16(h - r16(h))
        return r16(h) if self.qa else h

    def _grad(self, g):
        if self.split_a:
            return r16(g) + r16(g - r16(g))
        return r16(g) if self.qg else g

    def sdf_jacobian(self, code, xyz, max_batch=1 << 16):
        code = np.asarray(code, np.float32)
        x = np.asarray(xyz, np.float32)
        inp = np.concatenate([np.broadcast_to(code, (x.shape[0], code.shape[0])), x], axis=-1)
        L = self.code_len
        h = inp
        acts = []
        for k in range(self.n_layers):
            W = self.Wq[k] if 0 < k < self.n_layers - 1 else self.W[k]     # layer 0: hi/lo extras, layer 8: FP32
            if k in self.latent_in:
                # hidden part through the quantised weights, re-injected input exactly (hi/lo extras MMA)
                split = W.shape[1] - self.in_dim
                h = h @ W[:, :split].T + inp @ self.W[k][:, split:].T + self.b[k]
            else:
                h = h @ W.T + self.b[k]
            if k < self.n_layers - 1:
                acts.append(h > 0)
                h = np.maximum(h, 0)
                if k < self.n_layers - 2:
                    h = self._act(h)                                      # next layer's A operand
        out = np.tanh(h)
        g = (1 - out * out).astype(np.float32) @ self.W[self.n_layers - 1]
        g_in = 0
        for k in range(self.n_layers - 2, -1, -1):
            g = self._grad(g * acts[k])
            W = self.Wq[k]
            g = g @ W
            if k in self.latent_in:
                split = g.shape[1] - self.in_dim
                g_in = g_in + g[:, split:]
                g = g[:, :split]
        return out[:, 0], (g + g_in).astype(np.float32)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    sd = synth.load_wfit().state_dict()
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = 1
    variants = [("w", dict(qw=True)), ("a", dict(qa=True)), ("g", dict(qg=True)), ("w+a+g (tc mode)", dict(qw=True, qa=True, qg=True)),
                ("a+g, weights hi+lo", dict(split_w=True, qa=True, qg=True)),
                ("w, activations+gradients hi+lo", dict(qw=True, split_a=True))]
    dets = [synth.make_kitti_detection(k) for k in range(n)]
    ref = [orc.OracleOptimizer(orc.OracleDecoder(sd), cfg).reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
           for d in dets]
    for name, kw in variants:
        dc, ang = [], []
        for d, r in zip(dets, ref):
            o = orc.OracleOptimizer(QuantDecoder(sd, **kw), cfg).reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
            dc.append(np.linalg.norm(o["code"] - r["code"]) / np.linalg.norm(r["code"]))
            ang.append(synth.pose_errors(o["t_cam_obj"], r["t_cam_obj"])[1])
        print("%-34s code rel: median %.2e max %.2e | rot deg: median %.2e max %.2e"
              % (name, np.median(dc), np.max(dc), np.median(ang), np.max(ang)), flush=True)


if __name__ == "__main__":
    main()

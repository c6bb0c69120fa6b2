"""Make assets/wfit_decoder_fp16.npz: the seeded decoder fitted to the two-ellipsoid
car-like family of `dspslam.synth` (SURVEY.md §8(d) "W-fit").  Training is not
bit-reproducible across machines, so the result is committed; parameters are
rounded to fp16 so the file stays small and every consumer loads the exact same
fp32 values.   Usage:  python tools/make_wfit.py [steps]
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(ROOT, "dsp-slam-rgbd_b200"))
from dspslam import synth  # noqa: E402


def main(steps=1500, batch=4096, lr=5e-4):
    dec = synth.make_decoder(0)
    for p in dec.parameters():
        p.requires_grad_(True)
    opt = torch.optim.Adam(dec.parameters(), lr=lr)
    rng = np.random.default_rng(7)
    t0 = time.time()
    for it in range(steps):
        z = (0.3 * rng.normal(size=(batch, synth.CODE_LEN))).astype(np.float32)
        # 65 % of the samples near the surface, the rest uniform in the [-1,1] box
        axes, centres = synth.shape_params(z)
        prim = rng.integers(0, 2, size=batch)
        u = rng.normal(size=(batch, 3))
        u /= np.linalg.norm(u, axis=1, keepdims=True)
        ar = np.arange(batch)
        near = centres[ar, prim] + u * axes[ar, prim] * (1 + 0.12 * rng.normal(size=(batch, 1)))
        uni = rng.uniform(-1, 1, size=(batch, 3))
        x = np.where(rng.uniform(size=(batch, 1)) < 0.65, near, uni)
        target = synth.shape_sdf(x, axes, centres).astype(np.float32)
        inp = torch.from_numpy(np.concatenate([z, x.astype(np.float32)], 1))
        pred = dec(inp)[:, 0]
        loss = (torch.clamp(pred, -0.1, 0.1) - torch.from_numpy(target)).abs().mean()
        opt.zero_grad()
        loss.backward()
        opt.step()
        if it % 50 == 0 or it == steps - 1:
            print("step %4d  L1 %.5f  (%.0fs)" % (it, loss.item(), time.time() - t0), flush=True)
    os.makedirs(os.path.dirname(synth.WFIT_PATH), exist_ok=True)
    synth.state_dict_to_npz(dec.state_dict(), synth.WFIT_PATH)
    print("wrote", synth.WFIT_PATH, os.path.getsize(synth.WFIT_PATH), "bytes")


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 1500)

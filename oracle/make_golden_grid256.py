"""tests/golden/grid_sdf_g256_slab.npz: four x-planes (linear voxel indices [123 G^2, 127 G^2), 262 144 queries) of the
BASELINE C4 grid (G = 256) from the UNMODIFIED reference: its own `create_voxel_grid(256)` (the sheared grid of modern
torch) through its own `decode_sdf`, for the optimised code of C1 object 0.
TEST INFRASTRUCTURE; runs only in the build container.   Usage: python oracle/make_golden_grid256.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_goldens as mg  # noqa: E402

ref, synth, torch = mg.ref, mg.synth, mg.torch
G = 256
with np.load(synth.WFIT_PATH) as z:
    wfit = {k: torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
dec = mg.ref_decoder(wfit)
with np.load(os.path.join(mg.OUT, "grid_sdf_g64.npz")) as z:
    code = z["code"].astype(np.float32)
pts = ref.utils.create_voxel_grid(vol_dim=G)
i0, i1 = 123 * G * G, 127 * G * G
sdf = ref.loss_utils.decode_sdf(dec, torch.from_numpy(code), pts[i0:i1])
np.savez_compressed(os.path.join(mg.OUT, "grid_sdf_g256_slab.npz"), code=code, G=np.array(G), i0=np.array(i0),
                    i1=np.array(i1), sdf=sdf.numpy().astype(np.float32), xyz_first=pts[i0:i0 + 4].numpy())
print("done", sdf.shape, float(sdf.min()), float(sdf.max()))

"""tests/golden/bulk_c2.npz: the 16 detections of the benchmarked C2 batch itself (seeds 0..15, M = 2048, N = 2248,
D = 50, 10 iterations) through the UNMODIFIED reference: per-iteration state (T_oc, code), K, final pose / code / loss.
Inputs are regenerated from the seed (`dspslam.synth.make_kitti_detection(k, M=2048, n_bg=200)`); `in_sum` guards the
generator.  TEST INFRASTRUCTURE; runs only in the build container.   Usage: python oracle/make_goldens_c2bulk.py [n]"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_goldens as mg  # noqa: E402
import make_goldens_bulk as mb  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
with np.load(mg.synth.WFIT_PATH) as z:
    wfit = {k: mg.torch.from_numpy(z[k].astype(np.float32)) for k in z.files}
opt = mg.ref.optimizer.Optimizer(mg.ref_decoder(wfit), mg.ref_config("kitti"))
out = mb.bulk(opt, lambda k: mg.synth.make_kitti_detection(k, M=2048, n_bg=200), n, 10, False, keep_H=0)
out.pop("H", None)
np.savez_compressed(os.path.join(mg.OUT, "bulk_c2.npz"), **out)
print("done", out["is_good"].tolist())

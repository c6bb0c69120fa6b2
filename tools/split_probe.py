"""Experiment: the C2 batch as ONE launch sequence vs TWO half-batches on two streams (two library handles), so that
the small glue kernels and the wave tails of one half overlap the decoder kernels of the other."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402
from reconstruct.optimizer import Optimizer  # noqa: E402

dec = synth.load_wfit()
opt = Optimizer(dec, synth.kitti_config(), mode="tc")
prm = opt._params()
dets = [synth.make_kitti_detection(k, M=2048, n_bg=200) for k in range(16)]
nsplit = int(sys.argv[1]) if len(sys.argv) > 1 else 2
engs = [Engine(dec, mode="tc") for _ in range(nsplit)]
streams = [torch.cuda.Stream() for _ in range(nsplit)]
whole = opt.engine.upload_joint(opt.engine.pack_joint(dets))
parts = []
for i in range(nsplit):
    lo, hi = i * 16 // nsplit, (i + 1) * 16 // nsplit
    parts.append(engs[i].upload_joint(engs[i].pack_joint(dets[lo:hi])))
torch.cuda.synchronize()


def run_whole():
    opt.engine.run_joint(whole, prm)


def run_split():
    cur = torch.cuda.current_stream()
    for s in streams:
        s.wait_stream(cur)
    for e, s, p in zip(engs, streams, parts):
        with torch.cuda.stream(s):
            e.run_joint(p, prm)
    for s in streams:
        cur.wait_stream(s)


for name, fn in (("whole", run_whole), ("split%d" % nsplit, run_split), ("whole", run_whole), ("split%d" % nsplit, run_split)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print("%-7s %.3f ms per batch  %.1f objects/s" % (name, ms, 16 / ms * 1e3), flush=True)
ra = opt.engine.download_joint(whole)
rb = [e.download_joint(p) for e, p in zip(engs, parts)]
code = np.concatenate([r.code for r in rb])
print("bit-identical poses:", np.array_equal(ra.t_cam_obj, np.concatenate([r.t_cam_obj for r in rb])),
      "codes:", np.array_equal(ra.code, code))

"""Timing of the camera-object pose-graph kernels next to the numpy restatement.
A local-BA-sized graph (40 keyframes, 30 objects, 300 edges) and a global-BA-sized one (2 000 / 500 / 20 000)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT, os.path.join(ROOT, "tests")]
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402
from dspslam.posegraph import PoseGraph  # noqa: E402
from oracle import posegraph_oracle as pg  # noqa: E402  (timing comparison only)
from test_pose_graph import _random_graph  # noqa: E402

eng = Engine(synth.load_wfit())
for n_kf, n_obj, ne, cpu_edges in ((40, 30, 300, 300), (2000, 500, 20000, 1000)):
    verts, ei, ej, meas, _ = _random_graph(3, n_kf=n_kf, n_obj=n_obj, n_edges=ne)
    g = PoseGraph(eng, huber_delta=10.0)
    g.set_vertices(verts)
    g.set_edges(ei, ej, meas)
    for _ in range(3):
        g.linearize()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        g.linearize()
    b.record()
    torch.cuda.synchronize()
    t_gpu = a.elapsed_time(b) / 20 * 1e-3
    t0 = time.perf_counter()
    g.linearize().chi2.cpu()
    t_call = time.perf_counter() - t0
    t0 = time.perf_counter()
    pg.linearize(verts, ei[:cpu_edges], ej[:cpu_edges], meas[:cpu_edges], 1e3, 10.0)
    t_cpu = (time.perf_counter() - t0) * ne / cpu_edges
    print(json.dumps(dict(keyframes=n_kf, objects=n_obj, edges=ne, gpu_us_per_linearize=t_gpu * 1e6,
                          api_call_plus_readback_us=t_call * 1e6, numpy_restatement_ms=t_cpu * 1e3,
                          medges_per_s=ne / t_gpu / 1e6)), flush=True)

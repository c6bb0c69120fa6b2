"""Where does a layer step of the GN Jacobian kernel (k_tc<5>) spend its time?  Needs the trace build
(`make -C dsp-slam-rgbd_b200/csrc trace`); runs one C2 batch and prints, for CTA 0 of the last Jacobian launch, the
duration of every MMA phase (accumulator wait -> commit issued) and every epilogue (accumulator complete -> next A
operand signalled) in SM cycles.   python tools/trace_jac.py [mode]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
from dspslam import _native as N  # noqa: E402

N.LIB_PATH = os.path.join(os.path.dirname(N.LIB_PATH), "libdspslam_sm100_trace.so")
from dspslam import synth  # noqa: E402
from reconstruct.optimizer import Optimizer  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "tc"
opt = Optimizer(synth.load_wfit(), synth.kitti_config(), mode=mode)
eng = opt.engine
dets = [synth.make_kitti_detection(k, M=2048, n_bg=200) for k in range(16)]
batch = eng.upload_joint(eng.pack_joint(dets))
for _ in range(2):
    eng.run_joint(batch, opt._params())
buf = (C.c_longlong * 8192)()
eng.lib.dspslam_debug_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
assert eng.lib.dspslam_debug_trace(eng._h, buf, 8192) == 0
t = np.array(buf, dtype=np.int64)
ne, nm = int(t[4095]), int(t[8191])
E_ = t[:ne].reshape(-1, 2)
M_ = t[4096:4096 + nm].reshape(-1, 2)
t0 = min(E_[0, 0], M_[0, 0])
print("mode %s: %d epilogue events, %d MMA events of CTA 0" % (mode, len(E_), len(M_)))
# MMA phases: tag p*100+li = operands ready, 1000+... = commit issued (completion is seen by the epilogue's wait)
print("-- first tile, per layer step: [MMA issue window] and [epilogue: accumulator complete -> signalled]")
ev = sorted([(int(c - t0), "E", int(g)) for c, g in E_] + [(int(c - t0), "M", int(g)) for c, g in M_])
last = {}
rows = []
for c, w, g in ev[:140]:
    rows.append("%9d %s %5d" % (c, w, g))
print("\n".join(rows))
# summary over all tiles: epilogue durations by tag
dur = {}
start = {}
for c, g in E_:
    g = int(g)
    if g < 1000:
        start[g] = int(c)
    elif g < 2000 and (g - 1000) in start:
        dur.setdefault(g - 1000, []).append(int(c) - start[g - 1000])
print("-- epilogue durations (cycles), median over tiles: tag 0-6 forward layers, 7 = g7 build, 101-107 backward")
for g in sorted(dur):
    print("  %4d: %7.0f" % (g, np.median(dur[g])))
mstart, mdur = {}, {}
for c, g in M_:
    g = int(g)
    if g < 1000:
        mstart[g] = int(c)
    elif (g - 1000) in mstart:
        mdur.setdefault(g - 1000, []).append(int(c) - mstart[g - 1000])
print("-- MMA issue windows (cycles; operands ready -> last MMA of the layer issued), median over tiles")
for g in sorted(mdur):
    print("  %4d: %7.0f" % (g, np.median(mdur[g])))

# per-stage log of the first tile's forward layer 1: 5000+k before the wait for the stage, 6000+k stage present, 7000+k
# MMAs and commit issued
st = {}
for c, g in M_:
    g = int(g)
    if 5000 <= g < 8000:
        st.setdefault(g % 1000, {})[g // 1000] = int(c)
if st:
    print("-- forward layer 1, first tile: per weight stage [wait for the stage | issue 2 MMAs + commit | loop period]")
    prev = None
    for k in sorted(st):
        e = st[k]
        if 5 in e and 6 in e and 7 in e:
            print("  stage %2d: wait %5d  issue %5d  period %5s" % (k, e[6] - e[5], e[7] - e[6], "-" if prev is None else e[5] - prev))
            prev = e[5]

# last step of a tile: 100 accumulator complete, 1400 J row + weights written (warp 0), 1500 all rows there, 1600 warp 0's
# 4x4 blocks summed, 2000 partials stored and loss sums done
seq = [(int(c), int(g)) for c, g in E_ if int(g) in (100, 1400, 1500, 1550, 1600, 2000)]
rows, cur = [], {}
for c, g in seq:
    cur[g] = c
    if g == 2000 and len(cur) == 6:
        rows.append([cur[1400] - cur[100], cur[1500] - cur[1400], cur[1550] - cur[1500], cur[1600] - cur[1550], cur[2000] - cur[1600]])
        cur = {}
if rows:
    m = np.median(np.array(rows), axis=0)
    print("-- last step (cycles, median over tiles): J row + weights %d | wait for the other warps %d | J^T J row loop %d | partial "
          "stores %d | loss sums + barriers %d" % tuple(m))

"""The boundary exercised from C++ (Missing 4): tests/cpp_embed/embed_host.cpp embeds CPython through
pybind11, builds the objects on the main thread, releases the GIL and lets a "tracking" and a "local mapping" thread
re-take it around every call, passing column-major float matrices - the reference's own arrangement
(src/System.cc:91-99,158, include/System.h:57-71, src/LocalMapping_util.cc:165-206).  The CPU test covers the
interpreter / GIL hand-over / matrix round trip; the GPU test runs frame -> reconstruct_object -> mesh -> pose-only."""
import os
import subprocess
import sys
import sysconfig

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
PKG = os.path.join(ROOT, "dsp-slam-rgbd_b200")
EXE = os.path.join(ROOT, "tests", "cpp_embed", "embed_host")


def _build():
    import pybind11
    libdir, ver = sysconfig.get_config_var("LIBDIR"), sysconfig.get_config_var("LDVERSION")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-fvisibility=hidden", "-I", pybind11.get_include(), "-I",
           sysconfig.get_config_var("INCLUDEPY"), os.path.join(ROOT, "tests", "cpp_embed", "embed_host.cpp"), "-o", EXE,
           "-L", libdir, "-lpython" + ver, "-lpthread", "-ldl", "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def _env():
    # the embedded interpreter starts from the base installation: hand it this interpreter's module path (the venv)
    return dict(os.environ, PYTHONPATH=os.pathsep.join(p for p in sys.path if p))


def test_embedded_interpreter_gil_handover_and_matrix_round_trip():
    _build()
    r = subprocess.run([EXE, "--dry", PKG], capture_output=True, text=True, timeout=300, env=_env())
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-1500:])
    assert "embed host ok: detections=1 good=1" in r.stdout


@pytest.mark.gpu
def test_cpp_host_runs_the_reference_call_sequence(tmp_path):
    from test_boundary_replay import _scratch_kitti
    _build()
    seq_dir, cfg_path = _scratch_kitti(str(tmp_path))
    r = subprocess.run([EXE, PKG, seq_dir, cfg_path], capture_output=True, text=True, timeout=600, env=_env())
    print(r.stdout[-500:], r.stderr[-1500:])
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-1500:])
    line = [l for l in r.stdout.splitlines() if l.startswith("embed host ok")][-1]
    vals = dict(kv.split("=") for kv in line.split(": ")[1].split())
    assert int(vals["detections"]) >= 3 and int(vals["good"]) >= 2 and int(vals["faces"]) > 0

"""The C ABI from plain C: include/dspslam.h compiles as C (gcc -std=c11 -pedantic) and a C program linked against
libdspslam_sm100.so runs the decoder in every arithmetic mode without Python (tests/c_client/abi_client.c)."""
import os
import subprocess

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
LIBDIR = os.path.join(ROOT, "dsp-slam-rgbd_b200", "dspslam")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")
EXE = os.path.join(ROOT, "tests", "c_client", "abi_client")


def _build():
    cmd = ["gcc", "-std=c11", "-pedantic", "-Wall", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"), "-I",
           os.path.join(CUDA, "include"), os.path.join(ROOT, "tests", "c_client", "abi_client.c"), "-o", EXE,
           "-L", LIBDIR, "-l:libdspslam_sm100.so", "-L", os.path.join(CUDA, "lib64"), "-lcudart", "-lm",
           "-Wl,-rpath," + LIBDIR + ",-rpath," + os.path.join(CUDA, "lib64")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]


def test_header_is_c_and_client_links():
    _build()
    assert os.path.isfile(EXE)


@pytest.mark.gpu
def test_c_client_runs_every_mode():
    _build()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-500:])
    assert "c client ok" in r.stdout

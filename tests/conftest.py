import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
PKG = os.path.join(ROOT, "dsp-slam-rgbd_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session", autouse=True)
def _native_library_built():
    """A fresh checkout has no built library (built artefacts are git-ignored): build it once per session
    (`__graft_entry__.build()`: nvcc for sm_100a, works without a GPU)."""
    from dspslam import _native
    if not os.path.isfile(_native.LIB_PATH) or not os.path.isfile(os.path.join(PKG, "csrc", "libhd_selftest.so")):
        import __graft_entry__ as g
        g.build()

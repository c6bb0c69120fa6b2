"""dspslam_mean_sdf_noalloc against dspslam_mean_sdf, through the C ABI, on the GPU.

Some hosts of this library may not allocate device memory once they are running: a SLAM thread working inside a memory
budget fixed at start-up, a process whose GPU memory belongs to another allocator, a real-time loop that cannot afford
a driver allocation (it may block and synchronise the device).  For them the library has a variant of
dspslam_mean_sdf that parks the decoder outputs in a fixed 1 MiB array inside the library image instead of a workspace
grown with cudaMallocAsync.  This file checks that the variant returns what the allocating entry point returns.

The array is shared by every call and keeps the previous call's contents, so each case makes two calls in a row with
different inputs - a large one first, then a smaller one whose rows sit on top of the first call's leftovers - and
compares both.  The file is named to be collected last: `pytest -x` stops at the first failure, and a failure here
must not hide the parity tests.  It passes alone in a fresh process (the first call then does the per-GPU address
look-up) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from dspslam import synth
    from dspslam.engine import Engine
    return Engine(synth.load_wfit())


def _mean(eng, entry, mode, code, xyz):
    """One raw call of `entry` (a ctypes function of libdspslam_sm100.so) with device pointers; returns the float."""
    import torch
    from dspslam import _native as N
    out = torch.full((1,), float("nan"), dtype=torch.float32, device=eng.device)
    rc = entry(eng._h, eng._stream(), N.MODE_FP32 if mode == "fp32" else N.MODE_TC, code.data_ptr(), xyz.data_ptr(),
               xyz.shape[0], out.data_ptr())
    N.check(rc, entry.__name__)
    return float(out.cpu()[0])


@pytest.mark.parametrize("mode", ["fp32", "tc"])
def test_noalloc_matches_allocating_entry_point_twice_in_a_row(mode, eng):
    import torch
    lib = eng.lib
    rng = np.random.default_rng(1234)
    code = torch.from_numpy((0.05 * rng.standard_normal(eng.code_len)).astype(np.float32)).to(eng.device)
    tol = 1e-6 if mode == "fp32" else 1e-4          # the tolerances of test_objectpoint_loss
    for n in (5000, 777):                            # second call: fewer rows, on top of the first call's leftovers
        xyz = torch.from_numpy(rng.uniform(-0.9, 0.9, (n, 3)).astype(np.float32)).to(eng.device)
        ref = _mean(eng, lib.dspslam_mean_sdf, mode, code, xyz)
        got = _mean(eng, lib.dspslam_mean_sdf_noalloc, mode, code, xyz)
        assert np.isfinite(ref) and np.isfinite(got)
        assert abs(got - ref) <= tol, (mode, n, got, ref)


def test_noalloc_refuses_more_rows_than_the_array_holds(eng):
    """Beyond DSPSLAM_NOALLOC_MAX_ROWS the call must fail with a message, not write past the array.  The size check
    comes before any pointer is touched, so no buffer of that size is needed."""
    import torch
    from dspslam import _native as N
    lib = eng.lib
    code = torch.zeros(eng.code_len, dtype=torch.float32, device=eng.device)
    xyz = torch.zeros((8, 3), dtype=torch.float32, device=eng.device)
    out = torch.zeros(1, dtype=torch.float32, device=eng.device)
    rc = lib.dspslam_mean_sdf_noalloc(eng._h, eng._stream(), N.MODE_FP32, code.data_ptr(), xyz.data_ptr(),
                                      C.c_int64(N.NOALLOC_MAX_ROWS + 1), out.data_ptr())
    assert rc == 1                                   # DSPSLAM_E_BADARG
    assert b"DSPSLAM_NOALLOC_MAX_ROWS" in lib.dspslam_last_error()
    # at the limit itself the call goes through and agrees with the allocating entry point
    rng = np.random.default_rng(7)
    xyz = torch.from_numpy(rng.uniform(-0.9, 0.9, (N.NOALLOC_MAX_ROWS, 3)).astype(np.float32)).to(eng.device)
    ref = _mean(eng, lib.dspslam_mean_sdf, "fp32", code, xyz)
    got = _mean(eng, lib.dspslam_mean_sdf_noalloc, "fp32", code, xyz)
    assert abs(got - ref) <= 1e-6

"""CPU check of the host/device math header the kernels use (csrc/hd_math.h), compiled with g++
into a test-only shared object, against the oracle and the reference's golden vectors."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import dspslam_oracle as orc

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
CSRC = os.path.join(ROOT, "dsp-slam-rgbd_b200", "csrc")
G = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def hd(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("hd") / "libhd_selftest.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off",
                           os.path.join(CSRC, "hd_selftest.cpp"), "-o", so])
    L = C.CDLL(so)
    L.hd_huber.restype = C.c_float
    L.hd_huber.argtypes = [C.c_float, C.c_float]
    L.hd_rot_prior.restype = C.c_float
    L.hd_rot_prior.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
    L.hd_linspace_at.restype = C.c_float
    L.hd_linspace_at.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int]
    L.hd_ray_scan.restype = C.c_int
    L.hd_ray_scan.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p,
                              C.c_void_p, C.c_void_p]
    return L


def _call_mat(fn, x, n_out=16):
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(n_out, np.float32)
    fn(x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return out


def test_exp_maps_match_reference(hd):
    g = np.load(os.path.join(G, "liealg.npz"))
    for x, s7, s6 in zip(g["x"], g["exp_sim3"], g["exp_se3"]):
        assert np.abs(_call_mat(hd.hd_exp_sim3, x).reshape(4, 4) - s7).max() < 2e-6
        assert np.abs(_call_mat(hd.hd_exp_se3, x[:6]).reshape(4, 4) - s6).max() < 2e-6


def test_inverse_and_product(hd):
    rng = np.random.default_rng(0)
    for _ in range(10):
        T = np.eye(4, dtype=np.float32)
        T[:3, :3] = (np.linalg.qr(rng.normal(size=(3, 3)))[0] * rng.uniform(0.3, 3)).astype(np.float32)
        T[:3, 3] = rng.normal(scale=5, size=3)
        Ti = _call_mat(hd.hd_inv_affine, T).reshape(4, 4)
        assert np.abs(Ti - np.linalg.inv(T.astype(np.float64))).max() < 1e-5
        B = np.eye(4, dtype=np.float32)
        B[:3] = rng.normal(size=(3, 4))
        out = np.zeros(16, np.float32)
        hd.hd_mul_affine(T.ctypes.data_as(C.c_void_p), B.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert np.abs(out.reshape(4, 4) - T @ B).max() < 1e-5


def test_huber_and_rot_prior(hd):
    g = np.load(os.path.join(G, "liealg.npz"))
    for b in (0.025, 0.2, 0.05):
        got = np.array([hd.hd_huber(float(r), b) for r in g["hub_in"]], np.float32)
        assert np.abs(got - g["hub_r_%g" % b]).max() < 1e-7
    for T_oc, J, r in zip(g["rot_T_oc"], g["rot_J"], g["rot_res"]):
        T_co = np.linalg.inv(T_oc.astype(np.float64)).astype(np.float32)
        scale = np.float32(np.cbrt(np.linalg.det(T_co[:3, :3].astype(np.float64))))
        J3 = np.zeros(3, np.float32)
        res = hd.hd_rot_prior(np.ascontiguousarray(T_co).ctypes.data_as(C.c_void_p), float(scale),
                              J3.ctypes.data_as(C.c_void_p))
        assert abs(res - r) < 2e-6
        assert np.abs(J3 - J[3:6]).max() < 2e-5


def test_linspace(hd):
    a, b = np.float32(11.3), np.float32(15.9)
    ref = orc.linspace_torch(a, b, 50)
    got = np.array([hd.hd_linspace_at(float(a), float(b), 50, i) for i in range(50)], np.float32)
    assert np.array_equal(got, ref)
    import torch
    assert np.array_equal(torch.linspace(float(a), float(b), 50).numpy(), ref)


def test_ray_scan_against_oracle_render(hd):
    """One ray through hd_math's ray_scan == the oracle's vectorised render bookkeeping."""
    rng = np.random.default_rng(1)
    D, th = 50, 0.01
    for trial in range(200):
        dmin, dmax = np.float32(10.0), np.float32(14.0)
        # an SDF profile crossing zero somewhere, some samples outside the ball
        x = np.linspace(-1, 1, D)
        cross = rng.uniform(-0.6, 0.9)
        sdf = ((cross - x) * rng.uniform(0.01, 0.3) + rng.normal(scale=0.002, size=D)).astype(np.float32)
        outside = rng.uniform(size=D) < 0.2
        sdf_k = sdf.copy()
        sdf_k[outside] = np.float32(1e30)
        dobs = np.float32(rng.uniform(dmin, dmax))
        idx = np.zeros(D, np.int32); de = np.zeros(D, np.float32); rs = np.zeros(D, np.float32)
        n = hd.hd_ray_scan(sdf_k.ctypes.data_as(C.c_void_p), D, float(dobs), th, float(dmin), float(dmax),
                           idx.ctypes.data_as(C.c_void_p), de.ctypes.data_as(C.c_void_p), rs.ctypes.data_as(C.c_void_p))
        # oracle bookkeeping (loss.py:98-155) on the same profile
        f = np.float32
        occ = np.where(outside, f(0), f(0.5) - np.clip(sdf, f(-th), f(th)) / f(2 * th)).astype(f)
        wg = (~outside) & (sdf > -th) & (sdf < th)
        acc = np.cumprod(1 - occ, dtype=f)
        d = np.concatenate([orc.linspace_torch(dmin, dmax, D), [f(1.1) * dmax]]).astype(f)
        term = np.concatenate([occ, [f(1)]]) * np.concatenate([[f(1)], acc])
        d_u = (d * term).sum(dtype=f)
        suffix = np.cumsum(acc[::-1], dtype=f)[::-1]
        with np.errstate(divide="ignore", invalid="ignore"):
            de_do = suffix / (1 - occ)
        sel = np.nonzero(wg & (de_do > 1e-2))[0]
        assert n == sel.size and np.array_equal(idx[:n], sel)
        ref_de = de_do[sel] * ((dmax - dmin) / f(D - 1)) * f(-1 / (2 * th))
        assert np.allclose(de[:n], ref_de, rtol=2e-5, atol=1e-7)
        assert np.allclose(rs[:n], np.clip(dobs - d_u, -0.3, 0.3), atol=2e-5)

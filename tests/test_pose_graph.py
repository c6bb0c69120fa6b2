"""Camera-object pose-graph edges: include/ObjectPoseGraph.h:56-88 on g2o's SE3Quat.

PARITY UNPINNED: neither g2o nor Eigen can be built in this image and the reference has no vectors for this path, so
the CPU tests hold the restatement (oracle/posegraph_oracle.py) to the properties the arithmetic must have, and the
GPU tests hold the CUDA kernels (csrc/posegraph.cu, through the C ABI) to the restatement.  Tolerances: FP64 work,
|diff| <= 1e-12 * scale (the scale is 1 for poses / errors / Jacobians and inv_sigma = 1e3 for the quadratic form).
"""
import math
import os

import numpy as np
import pytest

from oracle import posegraph_oracle as pg

TOL = 1e-12


def _rand_pose(rng, rot=1.0, trans=5.0):
    return pg.se3_exp(np.concatenate([rng.normal(size=3) * rot, rng.normal(size=3) * trans]))


def _mat(p):
    T = np.eye(4)
    T[:3, :3] = pg.quat_to_matrix(p[3:])
    T[:3, 3] = p[:3]
    return T


# ---------------------------------------------------------------------------------------------------------------
# CPU: properties of the restatement
# ---------------------------------------------------------------------------------------------------------------
def test_log_exp_round_trip():
    rng = np.random.default_rng(0)
    for s in (0.02, 0.5, 1.5, 2.8):
        for _ in range(20):
            x = rng.normal(size=6)
            x[:3] *= s / np.linalg.norm(x[:3])
            y = np.array(pg.se3_log(pg.se3_exp(x)))
            assert np.abs(y - x).max() < 1e-11
    # g2o's small-angle branch (d > 0.99999) is first order in omega: exact only to O(theta^3)
    x = np.array([2e-3, -1e-3, 1.5e-3, 0.3, -0.2, 0.1])
    assert np.abs(np.array(pg.se3_log(pg.se3_exp(x))) - x).max() < 1e-8


def test_quaternion_matrix_round_trip_all_branches():
    rng = np.random.default_rng(1)
    axes = [np.array(a, float) for a in ((1, 0, 0), (0, 1, 0), (0, 0, 1))]
    cases = [rng.normal(size=3) for _ in range(20)]
    cases += [a * (math.pi - 1e-3) + rng.normal(size=3) * 1e-2 for a in axes]      # trace <= 0, each largest diagonal
    for w in cases:
        p = pg.se3_exp(np.concatenate([w, np.zeros(3)]))
        q = p[3:]
        assert q[3] >= 0 and abs(np.linalg.norm(q) - 1) < 1e-15
        R = np.array(pg.quat_to_matrix(q))
        assert np.abs(R @ R.T - np.eye(3)).max() < 1e-14
        q2 = pg.quat_normalized(pg.quat_from_matrix(R.tolist()))
        assert np.abs(np.array(q2) - np.array(q)).max() < 1e-14
        v = rng.normal(size=3)
        assert np.abs(np.array(pg.quat_rotate(q, v.tolist())) - R @ v).max() < 1e-14


def test_group_operations_match_homogeneous_matrices():
    rng = np.random.default_rng(2)
    for _ in range(20):
        a, b = _rand_pose(rng), _rand_pose(rng)
        assert np.abs(_mat(pg.se3_mul(a, b)) - _mat(a) @ _mat(b)).max() < 1e-13
        assert np.abs(_mat(pg.se3_inv(a)) - np.linalg.inv(_mat(a))).max() < 1e-13
        x = rng.normal(size=6) * 0.1
        lhs = pg.se3_log(pg.se3_mul(pg.se3_mul(a, pg.se3_exp(x)), pg.se3_inv(a)))
        assert np.abs(np.array(lhs) - pg.se3_adj(a) @ x).max() < 1e-12             # Ad(T) x = log(T exp(x) T^-1)


def test_from_matrix_is_converter_toSE3Quat():
    rng = np.random.default_rng(3)
    T = _mat(_rand_pose(rng)).astype(np.float32)
    p = pg.se3_from_matrix(T)
    assert np.abs(_mat(p) - T.astype(np.float64)).max() < 1e-6                     # FP32 rotation is only ~orthonormal
    assert p[:3] == [float(T[0, 3]), float(T[1, 3]), float(T[2, 3])]


def test_edge_error_zero_for_consistent_graph_and_jacobians_vs_differences():
    rng = np.random.default_rng(4)
    T_cw, T_ow = _rand_pose(rng), _rand_pose(rng)
    meas = pg.se3_mul(T_cw, pg.se3_inv(T_ow))                                      # T_co = T_cw T_wo
    assert np.abs(pg.edge_error(meas, T_cw, T_ow)).max() < 1e-12
    # a small residual: the analytic blocks are the first-order BCH form, exact up to O(|e|^2)
    T_ow2 = pg.se3_oplus(T_ow, rng.normal(size=6) * 0.01)
    e0 = np.array(pg.edge_error(meas, T_cw, T_ow2))
    Ji, Jj = pg.edge_jacobians(meas, e0.tolist())
    h = 1e-6
    for J, which in ((Ji, 0), (Jj, 1)):
        num = np.zeros((6, 6))
        for d in range(6):
            u = np.zeros(6)
            u[d] = h
            vp = (pg.se3_oplus(T_cw, u), T_ow2) if which == 0 else (T_cw, pg.se3_oplus(T_ow2, u))
            vm = (pg.se3_oplus(T_cw, -u), T_ow2) if which == 0 else (T_cw, pg.se3_oplus(T_ow2, -u))
            num[:, d] = (np.array(pg.edge_error(meas, *vp)) - np.array(pg.edge_error(meas, *vm))) / (2 * h)
        # Jj is the right-Jacobian inverse to first order: exact up to O(|e|^2).  Ji carries the same block where the
        # left-Jacobian inverse (opposite sign of the half-skew term) belongs, so it is off by O(|e|) |adj| - a
        # property of the reference's edge that the restatement keeps.
        bound = (np.abs(e0).max() * np.abs(J).max() if which == 0 else 2 * np.abs(e0).max() ** 2) + 1e-6
        assert np.abs(num - J).max() < bound, which
    # at zero residual they are exact: Ji = adj(Tij^-1), Jj = -I
    Ji0, Jj0 = pg.edge_jacobians(meas, [0.0] * 6)
    assert np.array_equal(Jj0, -np.eye(6)) and np.abs(Ji0 - pg.se3_adj(pg.se3_inv(meas))).max() == 0


def test_huber_kernel():
    d = 10.0
    assert pg.huber(99.0, d) == (99.0, 1.0) and pg.huber(100.0, d) == (100.0, 1.0)
    rho, w = pg.huber(400.0, d)
    assert rho == 2 * 20.0 * d - 100.0 and w == 0.5
    assert pg.huber(400.0, 0.0) == (400.0, 1.0)                                    # no kernel
    r1, _ = pg.huber(100.0 + 1e-9, d)
    assert abs(r1 - 100.0) < 1e-8                                                  # continuous at the threshold


def _random_graph(seed, n_kf=12, n_obj=7, n_edges=40, noise=0.02, outliers=2):
    rng = np.random.default_rng(seed)
    kf = [_rand_pose(rng, 0.5, 4.0) for _ in range(n_kf)]
    obj = [_rand_pose(rng, 1.5, 6.0) for _ in range(n_obj)]
    verts = kf + [pg.se3_oplus(o, rng.normal(size=6) * 0.05) for o in obj]          # perturbed object estimates
    ei = rng.integers(0, n_kf, n_edges)
    ej = rng.integers(0, n_obj, n_edges)
    ej[:n_obj] = np.arange(n_obj)                                                  # every object observed at least once
    meas = []
    for e in range(n_edges):
        m = pg.se3_mul(kf[ei[e]], pg.se3_inv(obj[ej[e]]))
        s = 1.0 if e >= outliers else 20.0
        meas.append(pg.se3_oplus(m, rng.normal(size=6) * noise * s))
    return np.array(verts), ei.astype(np.int32), (ej + n_kf).astype(np.int32), np.array(meas), np.array(kf + obj)


def test_quadratic_form_of_the_restatement():
    verts, ei, ej, meas, _ = _random_graph(5)
    o = pg.linearize(verts, ei, ej, meas, 1e3, 10.0)
    assert (o["w"] < 1).any() and (o["w"] == 1).any()                              # both Huber branches in the sample
    for v in range(len(verts)):
        assert np.abs(o["A"][v] - o["A"][v].T).max() <= 1e-9
        assert np.linalg.eigvalsh(o["A"][v]).min() > -1e-6
    e = 3
    W = o["w"][e] * 1e3
    assert np.allclose(o["Hij"][e], o["Ji"][e].T @ (W * o["Jj"][e]), rtol=0, atol=1e-9)
    assert np.isclose(o["chi2"][e], 1e3 * (o["err"][e] ** 2).sum(), rtol=1e-15)


# ---------------------------------------------------------------------------------------------------------------
# GPU: the CUDA kernels against the restatement, through the C ABI
# ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def graph_cls():
    from dspslam import synth
    from dspslam.engine import Engine
    from dspslam.posegraph import PoseGraph
    eng = Engine(synth.load_wfit())
    return lambda **kw: PoseGraph(eng, **kw)


@pytest.mark.gpu
def test_gpu_from_matrices(graph_cls):
    rng = np.random.default_rng(6)
    poses = [_rand_pose(rng) for _ in range(30)]
    poses += [pg.se3_exp(np.concatenate([np.eye(3)[k] * (math.pi - 1e-3), rng.normal(size=3)])) for k in range(3)]
    T = np.stack([_mat(p) for p in poses]).astype(np.float32)
    got = graph_cls().from_matrices(T).cpu().numpy()
    want = np.array([pg.se3_from_matrix(t) for t in T])
    assert got.shape == (33, 7) and np.abs(got - want).max() <= 4e-16
    assert graph_cls().from_matrices(np.zeros((0, 4, 4), np.float32)).shape == (0, 7)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,huber", [(7, 10.0), (8, math.sqrt(1e3)), (9, 0.0)])
def test_gpu_linearize_matches_restatement(graph_cls, seed, huber):
    verts, ei, ej, meas, _ = _random_graph(seed, n_kf=40, n_obj=25, n_edges=300)
    verts = np.concatenate([verts, verts[:2]])                                      # two vertices nobody refers to
    g = graph_cls(inv_sigma=1e3, huber_delta=huber)
    g.set_vertices(verts)
    g.set_edges(ei, ej, meas)
    lin = g.linearize()
    o = pg.linearize(verts, ei, ej, meas, 1e3, huber)
    for name, ref, scale in (("error", o["err"], 1.0), ("Ji", o["Ji"], 10.0), ("Jj", o["Jj"], 1.0),
                             ("chi2", o["chi2"], 1e3), ("rho", o["rho"], 1e3), ("weight", o["w"], 1.0),
                             ("A", o["A"], 1e5), ("b", o["b"], 1e4), ("Hij", o["Hij"], 1e4)):
        got = getattr(lin, name).cpu().numpy()
        assert got.shape == ref.shape, name
        assert np.abs(got - ref).max() <= TOL * scale * max(1.0, np.abs(ref).max() / scale), name
    assert np.array_equal(g.chi2().cpu().numpy(), lin.chi2.cpu().numpy())
    # vertices without edges get exact zeros
    lonely = np.setdiff1d(np.arange(len(verts)), np.concatenate([ei, ej]))
    assert lonely.size and not lin.A.cpu().numpy()[lonely].any() and not lin.b.cpu().numpy()[lonely].any()


@pytest.mark.gpu
def test_gpu_quadratic_form_across_edge_chunks(graph_cls):
    """5 000 edges: the per-vertex gather walks the edge list in chunks of 2 048 with one warp per half chunk; the
    sums must still be the edge-order sums of the restatement (and identical from run to run)."""
    verts, ei, ej, meas, _ = _random_graph(12, n_kf=60, n_obj=40, n_edges=5000)
    g = graph_cls(inv_sigma=1e3, huber_delta=10.0)
    g.set_vertices(verts)
    g.set_edges(ei, ej, meas)
    lin = g.linearize()
    o = pg.linearize(verts, ei, ej, meas, 1e3, 10.0)
    A, b = lin.A.cpu().numpy(), lin.b.cpu().numpy()
    assert np.abs(A - o["A"]).max() <= TOL * np.abs(o["A"]).max()
    assert np.abs(b - o["b"]).max() <= TOL * np.abs(o["b"]).max()
    again = g.linearize()
    assert np.array_equal(again.A.cpu().numpy(), A) and np.array_equal(again.b.cpu().numpy(), b)


@pytest.mark.gpu
def test_gpu_oplus_and_empty_graph(graph_cls):
    rng = np.random.default_rng(10)
    verts = np.array([_rand_pose(rng) for _ in range(17)])
    upd = rng.normal(size=(17, 6)) * 0.3
    upd[0] = 0.0
    upd[1, :3] = 1e-7                                                              # small-angle branch of exp
    g = graph_cls()
    g.set_vertices(verts)
    g.oplus(upd)
    want = np.array([pg.se3_oplus(list(v), u) for v, u in zip(verts, upd)])
    assert np.abs(g.vertices.cpu().numpy() - want).max() <= 1e-14
    g.set_edges(np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros((0, 7)))
    lin = g.linearize()
    assert lin.error.shape == (0, 6) and not lin.A.cpu().numpy().any() and lin.b.shape == (17, 6)


@pytest.mark.gpu
def test_gpu_bad_arguments_raise(graph_cls):
    from dspslam._native import NativeError
    g = graph_cls(inv_sigma=-1.0)
    g.set_vertices(np.array([pg.se3_exp(np.zeros(6))] * 2))
    g.set_edges([0], [1], [pg.se3_exp(np.zeros(6))])
    with pytest.raises(NativeError):
        g.linearize()
    with pytest.raises(ValueError):
        g.set_edges([0], [2], [pg.se3_exp(np.zeros(6))])
    with pytest.raises(ValueError):
        g.oplus(np.zeros((3, 6)))


@pytest.mark.gpu
def test_gpu_refine_objects_recovers_poses(graph_cls):
    """Objects against fixed keyframes: the robust cost never increases and the object poses return to the truth to
    within the measurement noise (rotation noise x lever arm) although two measurements are gross outliers."""
    verts, ei, ej, meas, truth = _random_graph(11, n_kf=20, n_obj=10, n_edges=120, noise=0.005, outliers=2)
    g = graph_cls(inv_sigma=1e3, huber_delta=10.0)
    g.set_vertices(verts)
    g.set_edges(ei, ej, meas)
    hist = g.refine_objects(np.arange(20, 30), iterations=6)
    assert all(b <= a * (1 + 1e-9) for a, b in zip(hist, hist[1:])) and hist[-1] < 0.8 * hist[0]
    out = g.vertices.cpu().numpy()
    assert np.array_equal(out[:20], verts[:20])                                    # keyframes untouched
    for v in range(20, 30):
        d = np.array(pg.se3_log(pg.se3_mul(list(out[v]), pg.se3_inv(list(truth[v])))))
        d0 = np.array(pg.se3_log(pg.se3_mul(list(verts[v]), pg.se3_inv(list(truth[v])))))
        assert np.linalg.norm(d) < 0.05 and np.linalg.norm(d) < 0.7 * np.linalg.norm(d0)
    # the outlier test the reference applies afterwards (chi2 > threshold) singles out the two bad measurements
    chi2 = g.chi2().cpu().numpy()
    assert set(np.argsort(chi2)[-2:]) == {0, 1}


@pytest.mark.gpu
def test_gpu_measurement_from_pose_only_optimiser(graph_cls):
    """Hand-over: the T_co that `estimate_pose_cam_obj` returns becomes the measurement of an edge without leaving
    float32/float64 conventions of the reference (Converter::toSE3Quat of a 4x4 FP32 matrix)."""
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    d = synth.make_kitti_detection(0)
    opt = Optimizer(synth.load_wfit(), synth.kitti_config(), mode="fp32")
    obj = opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
    assert obj.is_good
    T_co = np.asarray(obj.t_cam_obj, np.float32).copy()
    s = np.cbrt(np.linalg.det(T_co[:3, :3].astype(np.float64)))
    T_co[:3, :3] = (T_co[:3, :3] / s).astype(np.float32)                           # SE(3) part (MapObject::SE3Tco)
    g = graph_cls()
    T_cw = np.eye(4, dtype=np.float32)
    T_ow = np.linalg.inv(T_co.astype(np.float64)).astype(np.float32)               # object placed by this very detection
    v = g.from_matrices(np.stack([T_cw, T_ow]))
    g.set_vertices(v)
    g.set_edges([0], [1], g.from_matrices(T_co[None]))
    lin = g.linearize()
    assert float(lin.chi2[0]) < 1e-6 and np.abs(lin.error.cpu().numpy()).max() < 1e-4

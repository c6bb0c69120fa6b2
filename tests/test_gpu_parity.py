"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI
(dspslam.engine -> libdspslam_sm100.so) and through the reference-shaped Optimizer API, against
  * golden vectors produced by the reference's own code (tests/golden, oracle/make_goldens.py),
  * the numpy oracle on seeded inputs (ragged sizes, empty inputs, failure paths),
  * size-independent properties at the BASELINE sizes (grid slabs vs whole grid, batching invariance).
Tolerances (north_star): FP32 validation mode 1e-5 abs per SDF / Jacobian entry; tensor-core mode
1e-3 abs for the SDF and for >= 95 % of the Jacobian rows (every entry of the row; the others are rows with a
ReLU pre-activation within rounding of zero, see _jac_check); per-GN-step pose within 1 cm / 0.1 deg and code
within 1e-2 - per object here, over the 160-object population in tests/test_gpu_parity_bulk.py.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
G = os.path.join(ROOT, "tests", "golden")
MODES = ["fp32", "tc"]
TOL_SDF = {"fp32": 1e-5, "tc": 1e-3}


def load(name):
    with np.load(os.path.join(G, name)) as z:
        return {k: z[k] for k in z.files}


def params(cfg, iters=None):
    j = cfg.optimizer.joint_optim
    return dict(k1=j.k1, k2=j.k2, k3=j.k3, k4=j.k4, b1=j.b1, b2=j.b2, lr=j.learning_rate, s_damp=j.scale_damping,
                cut_off=cfg.optimizer.cut_off_threshold, num_iterations=j.num_iterations if iters is None else iters,
                num_depth_samples=cfg.optimizer.num_depth_samples)


@pytest.fixture(scope="module")
def wfit():
    from dspslam import synth
    return synth.load_wfit()


@pytest.fixture(scope="module")
def wrand():
    from dspslam import synth
    return synth.make_decoder(0)


@pytest.fixture(scope="module")
def eng(wfit):
    from dspslam.engine import Engine
    return Engine(wfit)


@pytest.fixture(scope="module")
def eng_rand(wrand):
    from dspslam.engine import Engine
    return Engine(wrand)


@pytest.fixture(scope="module")
def odec(wfit):
    from oracle import dspslam_oracle as orc
    return orc.OracleDecoder(wfit.state_dict())


def _jac_check(mode, jac, ref):
    err = np.abs(jac - ref)
    if mode == "fp32":
        # 1e-5 on every entry, except rows where a ReLU pre-activation lies within FP32 rounding of 0 and the two
        # summation orders land on different sides of it (measured: 5 rows in 38 011, 4 096 units per row): there the
        # derivative itself is discontinuous and single entries move by ~1e-3
        bad = err.max(axis=1) >= 1e-5
        assert bad.sum() <= jac.shape[0] // 4000 and err.max() < 5e-2
    else:
        # FP16 operands: 10-bit mantissa (same as TF32).  Entry errors are ~3e-4 rms on |J| ~ 0.5; a ReLU whose
        # pre-activation is within rounding of 0 flips and moves single entries by up to ~1e-1 (it does so between
        # any two FP32 implementations too, just more rarely).  Judge the distribution, not the max.
        assert np.sqrt((err ** 2).mean()) < 6e-4
        assert (err.max(axis=1) <= 1e-3).mean() >= 0.95
        assert (err <= 1e-3).mean() >= 0.998
        assert np.quantile(err, 0.999) < 4e-3


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("which", ["wfit", "wrand"])
def test_queries_vs_reference_golden(mode, which, eng, eng_rand):
    g = load("queries_%s.npz" % which)
    e = eng if which == "wfit" else eng_rand
    sdf = e.sdf(g["code"], g["xyz"], mode=mode).cpu().numpy()
    s2, jac = e.sdf_jac(g["code"], g["xyz"], mode=mode)
    assert np.abs(sdf - g["sdf"]).max() < TOL_SDF[mode]
    assert np.abs(s2.cpu().numpy() - g["sdf_j"]).max() < TOL_SDF[mode]
    _jac_check(mode, jac.cpu().numpy(), g["jac"])


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n", [1, 31, 127, 128, 129, 777, 5000, 38011])
def test_queries_ragged_sizes_vs_oracle(mode, n, eng, odec):
    # tile counts 1, 2, 7, 40 and 297: a lone tile (the pair's second CTA idles), odd counts below the grid size and
    # an odd count above it (148 persistent CTAs = 74 pairs, third trip with one idle CTA)
    rng = np.random.default_rng(n)
    code = (0.3 * rng.normal(size=64)).astype(np.float32)
    x = rng.uniform(-0.8, 0.8, size=(n, 3)).astype(np.float32)
    so, jo = odec.sdf_jacobian(code, x)
    assert np.abs(eng.sdf(code, x, mode=mode).cpu().numpy() - so).max() < TOL_SDF[mode]
    s, j = eng.sdf_jac(code, x, mode=mode)
    assert np.abs(s.cpu().numpy() - so).max() < TOL_SDF[mode]
    if n >= 777 or mode == "fp32":
        _jac_check(mode, j.cpu().numpy(), jo)


@pytest.mark.parametrize("n", [1, 63, 64, 65, 777, 9473, 38011])
def test_exact_mode_queries_ragged_sizes_vs_oracle(n, eng, odec):
    """DSPSLAM_MODE_TC_EXACT (64-row tiles of hi/lo stacked rows): tile counts 1, 2, 13, 149 (one more than the grid)
    and 594; SDF and every Jacobian entry at FP32 grade (typical row: largest entry error 2e-6) except rows with a ReLU
    unit at its kink: hi+lo FP16 pairs carry ~22 bits, so a unit within ~1e-6 of zero can still flip (measured 2 rows in
    777; the FP32 mode has 5 in 38 011)."""
    rng = np.random.default_rng(n + 1)
    code = (0.3 * rng.normal(size=64)).astype(np.float32)
    x = rng.uniform(-0.8, 0.8, size=(n, 3)).astype(np.float32)
    so, jo = odec.sdf_jacobian(code, x)
    s, j = eng.sdf_jac(code, x, mode="tc_exact")
    assert np.abs(s.cpu().numpy() - so).max() < 2e-5
    err = np.abs(j.cpu().numpy() - jo)
    assert (err.max(axis=1) > 1e-4).sum() <= max(2, n // 250), (err.max(axis=1) > 1e-4).sum()
    assert np.median(err.max(axis=1)) < 2e-5
    assert np.sqrt((err ** 2).mean()) < 5e-5 and err.max() < 0.2


def test_exact_mode_c2_sized_steps_vs_reference(eng):
    """One C2-sized detection through the exact mode, teacher-forced against the reference (H, b, next state)."""
    from dspslam import synth
    g = load("c2_obj1.npz")
    det = synth.make_kitti_detection(1, M=2048, n_bg=200)
    g.update(in_pts=det["pts"], in_rays=det["rays"], in_depth=det["depth"])
    _teacher_forced(eng, "tc_exact", g, synth.kitti_config(), range(10), 1e-2, 3e-2)


@pytest.mark.parametrize("mode", MODES)
def test_empty_query(mode, eng):
    code = np.zeros(64, np.float32)
    assert eng.sdf(code, np.zeros((0, 3), np.float32), mode=mode).shape[0] == 0


@pytest.mark.parametrize("mode", MODES)
def test_grid_vs_reference_golden_and_slabs(mode, eng, odec):
    from oracle import dspslam_oracle as orc
    g = load("grid_sdf.npz")
    Gd = int(g["G"])
    full = eng.grid_sdf(g["code"], Gd, mode=mode).cpu().numpy()
    assert np.abs(full.reshape(Gd, Gd, Gd) - g["sdf"]).max() < TOL_SDF[mode]
    # slabs reassemble to the whole grid bit-exactly (rows are independent)
    cuts = [0, 1000, 1777, 2048, Gd ** 3]
    parts = [eng.grid_sdf(g["code"], Gd, i0=a, i1=b, mode=mode).cpu().numpy() for a, b in zip(cuts[:-1], cuts[1:])]
    assert np.array_equal(np.concatenate(parts), full)
    lat = orc.grid_sdf(odec, g["code"], 12, "lattice")
    out = eng.grid_sdf(g["code"], 12, coord_mode="lattice", mode=mode).cpu().numpy().reshape(12, 12, 12)
    assert np.abs(out - lat).max() < TOL_SDF[mode]


def test_grid_full_size_properties(eng):
    """BASELINE C4 size (G=256): z-slab/x-slab split equals the single launch; TC vs FP32 agree on a sub-sample."""
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    Gd = 256
    full = eng.grid_sdf(code, Gd, mode="tc")
    half = Gd ** 3 // 2
    a = eng.grid_sdf(code, Gd, i0=0, i1=half, mode="tc")
    b = eng.grid_sdf(code, Gd, i0=half, i1=Gd ** 3, mode="tc")
    import torch
    assert torch.equal(torch.cat([a, b]), full)
    i0 = 123 * Gd * Gd
    ref = eng.grid_sdf(code, Gd, i0=i0, i1=i0 + 4 * Gd * Gd, mode="fp32")
    assert (full[i0:i0 + 4 * Gd * Gd] - ref).abs().max().item() < 1e-3
    assert full.min().item() < 0 < full.max().item()


def _teacher_forced(eng, mode, g, cfg, steps, tol_H, tol_b):
    """One GN step from the reference's own state at iteration e (pose, code), compared with the reference's H, b
    and next state.  FP32 mode: every step within the north_star tolerances.  Tensor-core mode: the Jacobian of a
    ReLU network is piecewise constant, and FP16-operand rounding flips a ReLU whose pre-activation is within ~1e-4
    of zero on a few percent of the rows; when such a row carries a large render coefficient one step can move H
    by a few percent (seen: 1 step of 20).  The population statement (>= 95 % of 1 280 + 640 steps and of 160 objects
    within the north_star tolerances) is asserted in tests/test_gpu_parity_bulk.py; with the 10 - 20 steps of ONE
    object a single miss is already 5 - 10 %, so here tensor-core mode asserts at most 10 % misses at the same
    tolerances as the FP32 mode and every step within loose bounds."""
    from dspslam import synth
    p = params(cfg, iters=1)
    n_it = g["H"].shape[0]
    strict_ok = []
    code_tol = 1e-2
    for e in steps:
        T_co = np.linalg.inv(g["T_oc"][e].astype(np.float64)).astype(np.float32)
        res = eng.gn_joint([(T_co, g["in_pts"], g["in_rays"], g["in_depth"], g["code_it"][e])], p, mode=mode, debug=True)
        d = res.dbg[0, 0]
        H, b = d[:5041].reshape(71, 71), d[5041:5112]
        assert res.status[0] == 0
        # the reference's own row count moves by 1-2 rows in ~400 between 1 and 8 CPU threads
        Kref = int(g["K"][e])
        assert abs(int(d[5112]) - Kref) <= (max(2, Kref // 300) if mode == "fp32" else max(4, Kref // 100)), (e, d[5112], Kref)
        eH = np.linalg.norm(H - g["H"][e]) / np.linalg.norm(g["H"][e])
        eb = np.linalg.norm(b - g["b"][e]) / np.linalg.norm(g["b"][e])
        Tn, cn = ((np.linalg.inv(g["T_oc"][e + 1].astype(np.float64)), g["code_it"][e + 1]) if e + 1 < n_it
                  else (g["t_cam_obj"], g["code"]))
        dt, ang, _ = synth.pose_errors(res.t_cam_obj[0], Tn)
        ec = np.linalg.norm(res.code[0] - cn) / max(np.linalg.norm(cn), 0.1)
        ok = eH < tol_H and eb < tol_b and dt < 1e-2 and ang < 0.1 and ec < code_tol     # north_star: 1 cm / 0.1 deg
        strict_ok.append(ok)
        if mode == "fp32":
            assert ok, (e, eH, eb, dt, ang, ec)
        else:
            assert eH < 0.15 and eb < 0.15 and dt < 1e-2 and ang < 0.5 and ec < 0.1, (e, eH, eb, dt, ang, ec)
    assert np.mean(strict_ok) >= 0.9, strict_ok


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("k", [0, 1, 2, 3])
def test_c1_gn_steps_vs_reference(mode, k, eng):
    from dspslam import synth
    tol = (3e-3, 1e-2) if mode == "fp32" else (3e-2, 6e-2)
    _teacher_forced(eng, mode, load("c1_obj%d.npz" % k), synth.kitti_config(), range(10), *tol)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("k", [0, 1])
def test_c3_mono_gn_steps_vs_reference(mode, k, eng):
    from dspslam import synth
    # render rows dominate here (k1 = 10): a handful of rows entering/leaving the |s|<th, de_do>1e-2 selection
    # (K differs by a few of ~400) moves H by a few percent in tensor-core mode
    tol = (3e-3, 1e-2) if mode == "fp32" else (6e-2, 8e-2)
    steps = [0, 1, 5, 10, 15, 19] if mode == "fp32" else range(20)
    _teacher_forced(eng, mode, load("c3_obj%d.npz" % k), synth.redwood_config(), steps, *tol)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("k", [0, 1])
def test_c3_freiburg_gn_steps_vs_reference(mode, k, eng):
    """The third dataset configuration of the reference (configs/config_freiburg_001.json: monocular cars, k1 = 1,
    k3 = 0.5, scale damping 100, 50 depth samples), vectors from the reference's own run (make_goldens_freiburg.py)."""
    from dspslam import synth
    tol = (3e-3, 1e-2) if mode == "fp32" else (6e-2, 8e-2)
    _teacher_forced(eng, mode, load("c3f_obj%d.npz" % k), synth.freiburg_config(iters=10, num_depth_samples=50),
                    range(10), *tol)


@pytest.mark.parametrize("mode", MODES)
def test_reconstruct_object_api_free_run(mode, wfit):
    """Through the reference-shaped API.  The free-running trajectory is chaotic in the reference itself:
    oracle/noise_floor.py (tests/golden/noise_floor.json) shows the UNMODIFIED reference ending up to 0.09 deg /
    5e-2 relative code away from itself when the surface points move by one FP32 ulp, and 0.04 deg / 1.4e-2 with
    1 instead of 8 CPU threads.  So the end state only has to land near the reference's; per-step parity at the
    north_star tolerances is asserted by the teacher-forced tests above."""
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    opt = Optimizer(wfit, synth.kitti_config(), mode=mode)
    assert opt.code_len == 64
    n_ok = 0
    for k in range(4):
        g = load("c1_obj%d.npz" % k)
        T_in = g["in_t_cam_obj"].copy()
        out = opt.reconstruct_object(np.asfortranarray(T_in), g["in_pts"], g["in_rays"], g["in_depth"])
        assert np.array_equal(T_in, g["in_t_cam_obj"])                 # caller arrays untouched
        assert out.is_good and out.t_cam_obj.shape == (4, 4) and out.t_cam_obj.dtype == np.float32
        assert out.code.shape == (64,) and out.code.dtype == np.float32 and isinstance(out.loss, float)
        dt, ang, _ = synth.pose_errors(out.t_cam_obj, g["t_cam_obj"])
        # FP32 mode wanders by the reference's own noise floor (<= 0.1 deg here).  The tensor-core mode's 1e-4 SDF
        # error is 1000x that perturbation: over 64 detections its end state sits 0.14 deg (median) / 1.4 deg (max)
        # from the FP32 mode's (tools/free_run_stats.py) at the same final loss - the valley is flat along yaw/code.
        if mode == "fp32":
            assert dt < 0.01 and ang < 0.2 and np.linalg.norm(out.code - g["code"]) < 0.01
        else:
            assert dt < 0.02 and ang < 2.0 and np.linalg.norm(out.code - g["code"]) < 0.05
        assert abs(out.loss - float(g["loss"])) < 2e-2
        n_ok += (ang < 0.1)
    assert n_ok >= (4 if mode == "fp32" else 1)


@pytest.mark.parametrize("mode", MODES)
def test_batching_invariance(mode, eng):
    """B detections in one batch == the same detections one by one (objects are independent)."""
    from dspslam import synth
    p = params(synth.kitti_config(), iters=3)
    dets = [synth.make_kitti_detection(k, M=100 + 17 * k, n_bg=40 + 5 * k) for k in range(5)]
    batch = eng.gn_joint(dets, p, mode=mode)
    for k, d in enumerate(dets):
        one = eng.gn_joint([d], p, mode=mode)
        assert one.status[0] == batch.status[k]
        assert np.allclose(one.t_cam_obj[0], batch.t_cam_obj[k], atol=1e-5)
        assert np.allclose(one.code[0], batch.code[k], atol=1e-5)


@pytest.mark.parametrize("mode", MODES)
def test_failure_paths(mode, wrand, wfit, eng):
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    g = load("wrand_fail.npz")
    out = Optimizer(wrand, synth.kitti_config(), mode=mode).reconstruct_object(
        g["in_t_cam_obj"], g["in_pts"], g["in_rays"], g["in_depth"])
    assert not out.is_good and out.t_cam_obj is None and out.code is None      # reference: is_good False
    # too few ray samples inside the unit ball (rays looking away) -> is_good False; good object next to it unaffected
    d = synth.make_kitti_detection(1)
    bad = dict(d, rays=d["rays"] * np.array([-1, 1, 1], np.float32) + np.array([3, 0, 0], np.float32))
    res = eng.gn_joint([bad, d], params(synth.kitti_config(), iters=2), mode=mode)
    assert res.status[0] == 1 and res.status[1] == 0
    # empty rays
    empty = dict(d, rays=np.zeros((0, 3), np.float32), depth=np.zeros((0,), np.float32))
    res = eng.gn_joint([empty], params(synth.kitti_config(), iters=1), mode=mode)
    assert res.status[0] == 1


@pytest.mark.parametrize("mode", MODES)
def test_pose_only_vs_reference(mode, wfit):
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    opt = Optimizer(wfit, synth.kitti_config(), mode=mode)
    for k in range(2):
        g = load("pose_only_obj%d.npz" % k)
        T_in = g["in_t_co_se3"].copy()
        T = opt.estimate_pose_cam_obj(T_in, float(g["in_scale"]), g["in_pts"], g["in_code"])
        assert np.array_equal(T_in, g["in_t_co_se3"]) and T.shape == (4, 4)
        assert np.abs(T - g["t_cam_obj"]).max() < (1e-4 if mode == "fp32" else 5e-3)
        _, dbg = opt.estimate_poses_cam_obj([(T_in, float(g["in_scale"]), g["in_pts"], g["in_code"])], debug=True)
        H0, b0 = dbg[0, 0, :36].reshape(6, 6), dbg[0, 0, 36:42]
        tol = 1e-3 if mode == "fp32" else 2e-2
        assert np.linalg.norm(H0 - g["H"][0]) / np.linalg.norm(g["H"][0]) < tol
        assert np.linalg.norm(b0 - g["b"][0]) / np.linalg.norm(g["b"][0]) < tol


def test_pose_only_inlier_cut_vs_oracle(wfit, odec):
    """num_iterations > 5 exercises the e == 4 inlier cut (optimizer.py:77-79)."""
    from dspslam import synth
    from oracle import dspslam_oracle as orc
    from reconstruct.optimizer import Optimizer
    cfg = synth.kitti_config(pose_only_iters=8)
    g = load("pose_only_obj0.npz")
    pts = g["in_pts"].copy()
    pts[::7] += 0.4                                                     # outliers
    ref = orc.OracleOptimizer(odec, cfg).estimate_pose_cam_obj(g["in_t_co_se3"].copy(), float(g["in_scale"]), pts, g["in_code"])
    T = Optimizer(wfit, cfg, mode="fp32").estimate_pose_cam_obj(g["in_t_co_se3"], float(g["in_scale"]), pts, g["in_code"])
    assert np.abs(T - ref).max() < 2e-3


@pytest.mark.parametrize("mode", MODES)
def test_objectpoint_loss(mode, wfit):
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    g = load("objectpoint.npz")
    v = Optimizer(wfit, synth.kitti_config(), mode=mode).compute_sdf_loss_objectpoint_zhjd(g["pts"], g["code"])
    assert abs(v - float(g["mean_sdf"])) < (1e-6 if mode == "fp32" else 1e-4)


def test_c2_sized_batch_tc_vs_fp32(eng):
    """BASELINE C2 shape (2k surface points, 2k+200 rays, 50 samples) - one GN step, tensor-core vs FP32 mode."""
    from dspslam import synth
    p = params(synth.kitti_config(), iters=1)
    dets = [synth.make_kitti_detection(k, M=2048, n_bg=200) for k in range(3)]
    a = eng.gn_joint(dets, p, mode="tc", debug=True)
    b = eng.gn_joint(dets, p, mode="fp32", debug=True)
    for k in range(3):
        assert a.status[k] == 0 and b.status[k] == 0
        Ha, Hb = a.dbg[0, k, :5041], b.dbg[0, k, :5041]
        assert np.linalg.norm(Ha - Hb) / np.linalg.norm(Hb) < 3e-2
        dt, ang, _ = synth.pose_errors(a.t_cam_obj[k], b.t_cam_obj[k])
        assert dt < 1e-2 and ang < 0.1


def test_c2_sized_object_fp32_vs_oracle(eng, odec):
    """One GN step of a C2-sized detection (2 048 surface points + a few thousand render rows = ~40 row groups, so the
    partial reduction of the solve kernel takes its 16-wide path) against the oracle's H, b and updated pose."""
    from oracle import dspslam_oracle as orc
    from dspslam import synth
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = 1
    d = synth.make_kitti_detection(1, M=2048, n_bg=200)
    res = eng.gn_joint([d], params(cfg, iters=1), mode="fp32", debug=True)
    trace = []
    o = orc.OracleOptimizer(odec, cfg).reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"], trace=trace)
    dbg = res.dbg[0, 0]
    H, b = dbg[:5041].reshape(71, 71), dbg[5041:5112]
    assert res.status[0] == 0 and o["is_good"]
    assert abs(int(dbg[5112]) - int(trace[0]["K"])) <= 2 and int(trace[0]["K"]) + 2048 > 16 * 128
    assert np.linalg.norm(H - trace[0]["H"]) / np.linalg.norm(trace[0]["H"]) < 1e-3
    assert np.linalg.norm(b - trace[0]["b"]) / np.linalg.norm(trace[0]["b"]) < 1e-3
    dt, ang, _ = synth.pose_errors(res.t_cam_obj[0], o["t_cam_obj"])
    assert dt < 1e-3 and ang < 1e-2


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("k", [0, 1, 2, 3])
def test_c2_sized_gn_steps_vs_reference(mode, k, eng):
    """BASELINE C2 size against the reference itself: M = 2048 surface points, 2248 rays x 50 samples, ~2.5 k render
    rows per iteration; per-iteration state, H, b and K from the unmodified reference (oracle/make_goldens_bulk.py,
    ~10 s per object on CPU).  The inputs are regenerated from the seed and checked against the stored checksum."""
    from dspslam import synth
    g = load("c2_obj%d.npz" % k)
    det = synth.make_kitti_detection(k, M=2048, n_bg=200)
    chk = float(sum(np.asarray(det[n], np.float64).sum() for n in ("t_cam_obj", "pts", "rays", "depth")))
    assert abs(chk - float(g["in_sum"])) < 1e-6 * max(1.0, abs(chk))
    g.update(in_pts=det["pts"], in_rays=det["rays"], in_depth=det["depth"])
    tol = (3e-3, 1e-2) if mode == "fp32" else (3e-2, 6e-2)
    _teacher_forced(eng, mode, g, synth.kitti_config(), range(10), *tol)


@pytest.mark.parametrize("mode", MODES)
def test_grid_g64_vs_reference_golden(mode, eng):
    """MeshExtractor's grid (optimizer.py:224-228) at G = 64 for an optimised code, from the reference's own
    `decode_sdf` over its sheared `create_voxel_grid`."""
    g = load("grid_sdf_g64.npz")
    out = eng.grid_sdf(g["code"], 64, mode=mode).cpu().numpy().reshape(64, 64, 64)
    assert np.abs(out - g["sdf"]).max() < TOL_SDF[mode]


@pytest.mark.parametrize("mode", MODES)
def test_grid_g256_slab_vs_reference_golden(mode, eng):
    """BASELINE C4 size against the reference itself: four x-planes (262 144 voxels) of the G = 256 grid, the
    reference's own sheared `create_voxel_grid(256)` through its own `decode_sdf` (oracle/make_golden_grid256.py)."""
    g = load("grid_sdf_g256_slab.npz")
    G_, i0, i1 = int(g["G"]), int(g["i0"]), int(g["i1"])
    out = eng.grid_sdf(g["code"], G_, i0=i0, i1=i1, mode=mode).cpu().numpy()
    assert out.shape == g["sdf"].shape
    assert np.abs(out - g["sdf"]).max() < TOL_SDF[mode]
    assert g["sdf"].min() < 0 < g["sdf"].max()


def test_grid_sharded_equals_single(eng):
    """dspslam.dist.grid_sdf_sharded (world size 1 here; tools/dist_check.py runs it under torchrun + NCCL)."""
    import torch
    from dspslam.dist import grid_sdf_sharded
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    a = grid_sdf_sharded(eng, code, 24, mode="tc")
    b = eng.grid_sdf(code, 24, mode="tc").view(24, 24, 24)
    assert torch.equal(a, b)


@pytest.mark.parametrize("mode", MODES)
def test_ray_phases_bit_identical(mode, eng):
    """Early ray termination (dspslam_joint_params.ray_phases): a ray is not sampled behind its first sample with
    sdf <= -th, where occupancy is exactly 1 and the transmittance exactly 0 (loss.py:98-139).  Every phase count
    must give bit-identical H, b, poses and codes, while the decoder evaluates fewer rows."""
    from dspslam import synth
    dets = [synth.make_kitti_detection(k, M=300 + 40 * k, n_bg=80 + 9 * k) for k in range(4)]
    base, rows = None, {}
    for ph in (1, 2, 3, 4):
        p = dict(params(synth.kitti_config(), iters=4), ray_phases=ph)
        eng.profile_enable(True)
        r = eng.gn_joint(dets, p, mode=mode, debug=True)
        prof = eng.profile_read()
        eng.profile_enable(False)
        rows[ph] = (prof["rows_gn_fwd"], prof["rows_gn_inball"])
        assert (r.status == 0).all()
        if base is None:
            base = r
            assert prof["rows_gn_fwd"] == prof["rows_gn_inball"] > 0      # one pass evaluates every in-ball sample
            continue
        assert np.array_equal(r.status, base.status)
        assert np.array_equal(r.dbg, base.dbg)
        assert np.array_equal(r.t_cam_obj, base.t_cam_obj) and np.array_equal(r.code, base.code)
        assert np.array_equal(r.loss, base.loss)
        assert prof["rows_gn_inball"] == rows[1][1]
        assert prof["rows_gn_fwd"] < 0.8 * rows[1][0]


def test_flip_hypotheses_batched_vs_oracle(wfit, odec):
    """Mono orientation disambiguation (src/LocalMapping_util.cc:398-410): the pose and its twin with object x/z
    negated are optimised in ONE batch; the kept result is the first unless its loss is strictly greater.  FP32 mode
    against the oracle running the two hypotheses one after the other, as the reference's C++ does."""
    from dspslam import synth
    from oracle import dspslam_oracle as orc
    from reconstruct.optimizer import Optimizer
    cfg = synth.redwood_config(iters=3)
    opt = Optimizer(wfit, cfg, mode="fp32")
    oopt = orc.OracleOptimizer(odec, cfg)
    dets = [synth.make_mono_detection(k, M=96, n_fg=160, n_bg=60) for k in range(3)]
    # start the middle one from the wrong half turn so that the flipped twin is the one to keep
    dets[1] = dict(dets[1], t_cam_obj=Optimizer.flipped(dets[1]["t_cam_obj"]))
    both = opt.reconstruct_objects_flip(dets, return_both=True)
    kept = opt.reconstruct_objects_flip(dets)
    for d, (a, b), k in zip(dets, both, kept):
        ra = oopt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"], d["code"])
        rb = oopt.reconstruct_object(Optimizer.flipped(d["t_cam_obj"]), d["pts"], d["rays"], d["depth"], d["code"])
        for out, ref in ((a, ra), (b, rb)):
            assert out.is_good == bool(ref["is_good"])
            # three free-running iterations: the losing hypothesis sits in the wrong basin, where the row selections
            # make the trajectory discontinuous, so it is only compared on its loss
            assert abs(out.loss - float(ref["loss"])) < 1e-2
        win, rwin = (b, rb) if a.loss > b.loss else (a, ra)
        assert (float(ra["loss"]) > float(rb["loss"])) == (a.loss > b.loss)     # same hypothesis kept as the oracle
        if win.is_good:
            dt, ang, _ = synth.pose_errors(win.t_cam_obj, rwin["t_cam_obj"])
            assert dt < 1e-2 and ang < 0.2 and np.linalg.norm(win.code - rwin["code"]) < 2e-2
        want = b if a.loss > b.loss else a
        assert k.loss == want.loss and (k.t_cam_obj is None or np.array_equal(k.t_cam_obj, want.t_cam_obj))
        # a single two-hypothesis call gives the same answer as the batch
    one = opt.reconstruct_object_flip(dets[1]["t_cam_obj"], dets[1]["pts"], dets[1]["rays"], dets[1]["depth"], dets[1]["code"])
    assert abs(one.loss - kept[1].loss) < 1e-5
    # the wrong half turn must lose: the kept hypothesis of detection 1 is the flipped twin
    assert both[1][1].loss < both[1][0].loss and kept[1].loss == both[1][1].loss


def test_two_host_threads_share_one_optimizer(wfit):
    """The boundary's threading model: the tracking and local-mapping threads of the C++ system call
    the same Optimizer; ctypes releases the GIL during the native call, so calls do overlap here.  Results must
    equal the serial ones bit for bit (per-handle mutex, per-call workspace carve-up)."""
    import threading
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = 3
    opt = Optimizer(wfit, cfg, mode="tc")
    dets = [synth.make_kitti_detection(k, M=150 + 30 * k, n_bg=50 + 10 * k) for k in range(6)]
    serial = [opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"]) for d in dets]
    got = [None] * len(dets)
    errs = []

    def work(lo, hi):
        try:
            for _ in range(3):
                for k in range(lo, hi):
                    d = dets[k]
                    got[k] = opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
        except Exception as e:                                    # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=work, args=(0, 3)), threading.Thread(target=work, args=(3, 6))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs
    for a, b in zip(serial, got):
        assert a.is_good == b.is_good and a.loss == b.loss
        assert np.array_equal(a.t_cam_obj, b.t_cam_obj) and np.array_equal(a.code, b.code)


def test_degenerate_inputs_do_not_raise(wfit):
    """No exception may cross the pybind11 boundary: empty batches, missing points, NaN inputs and other
    dtypes / memory orders give the reference's answers (is_good False / NaN) instead of raising."""
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = 2
    opt = Optimizer(wfit, cfg, mode="tc")
    d = synth.make_kitti_detection(0)
    base = opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
    assert base.is_good
    f64 = opt.reconstruct_object(d["t_cam_obj"].astype(np.float64), np.asfortranarray(d["pts"].astype(np.float64)),
                                 d["rays"].astype(np.float64), d["depth"].astype(np.float64))
    assert f64.is_good and np.array_equal(f64.t_cam_obj, base.t_cam_obj) and np.array_equal(f64.code, base.code)
    assert opt.reconstruct_objects([]) == []
    assert not opt.reconstruct_object(d["t_cam_obj"], np.zeros((0, 3), np.float32), d["rays"], d["depth"]).is_good
    assert not opt.reconstruct_object(d["t_cam_obj"], np.full_like(d["pts"], np.nan), d["rays"], d["depth"]).is_good
    assert not opt.reconstruct_object(np.full((4, 4), np.nan, np.float32), d["pts"], d["rays"], d["depth"]).is_good
    long_code = opt.reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"], np.zeros(100, np.float32))
    assert long_code.is_good and long_code.code.shape == (64,)
    T = np.eye(4, dtype=np.float32)
    T[2, 3] = 10.0
    out = opt.estimate_poses_cam_obj([(T, 2.0, np.zeros((0, 3), np.float32), np.zeros(64, np.float32)),
                                      (d["t_cam_obj"] / np.float32(1), 2.0, d["pts"], np.zeros(64, np.float32))])
    assert out.shape == (2, 4, 4) and np.isnan(out[0]).all() and np.isfinite(out[1]).all()
    assert opt.estimate_poses_cam_obj([]).shape == (0, 4, 4)
    assert np.isnan(opt.compute_sdf_loss_objectpoint_zhjd(np.zeros((0, 3), np.float32), np.zeros(64, np.float32)))


@pytest.mark.parametrize("mode", ["fp32", "tc_exact", "tc"])
def test_gn_fuzz_ragged_shapes_vs_oracle(mode, eng, odec):
    """Seeded fuzz of the joint driver against the oracle: ragged surface / ray counts (down to a handful), depth-sample
    counts from 16 to 64, hyper-parameters of all three configurations, warm-start codes, 2 GN iterations, as ONE
    ragged batch.  FP32 and exact modes: pose within 1 mm / 0.02 deg of the oracle per object; default mode: the
    north_star per-step tolerances on most objects."""
    from oracle import dspslam_oracle as orc
    from dspslam import synth
    rng = np.random.default_rng(2026)
    cases = []
    for k in range(7):
        D = int(rng.choice([16, 23, 32, 48, 50, 64]))
        M = int(rng.integers(12, 700))
        n_bg = int(rng.integers(0, 150))
        kind = int(rng.integers(0, 3))
        cases.append((k, D, M, n_bg, kind))
    by_cfg = {}
    for k, D, M, n_bg, kind in cases:
        by_cfg.setdefault((D, kind), []).append((k, M, n_bg))
    n_ok, n_all = 0, 0
    for (D, kind), items in by_cfg.items():
        cfg = [synth.kitti_config, synth.redwood_config, synth.freiburg_config][kind]()
        cfg.optimizer.num_depth_samples = D
        cfg.optimizer.joint_optim.num_iterations = 2
        dets = []
        for k, M, n_bg in items:
            d = synth.make_kitti_detection(300 + k, M=M, n_bg=n_bg)
            d["code"] = (0.05 * np.random.default_rng(k).normal(size=64)).astype(np.float32)
            dets.append(d)
        res = eng.gn_joint(dets, params(cfg, iters=2), mode=mode)
        for i, d in enumerate(dets):
            ref = orc.OracleOptimizer(odec, cfg).reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"], d["depth"], code=d["code"])
            assert bool(ref["is_good"]) == (res.status[i] == 0), (D, kind, items[i])
            if not ref["is_good"]:
                continue
            dt, ang, _ = synth.pose_errors(res.t_cam_obj[i], ref["t_cam_obj"])
            ec = np.linalg.norm(res.code[i] - ref["code"]) / max(np.linalg.norm(ref["code"]), 0.1)
            n_all += 1
            if mode == "tc":
                n_ok += (dt < 1e-2 and ang < 0.1 and ec < 1e-2)
                assert dt < 2e-2 and ang < 0.5 and ec < 0.1, (mode, D, kind, items[i], dt, ang, ec)
            else:
                n_ok += 1
                assert dt < 1e-3 and ang < 0.02 and ec < 5e-3, (mode, D, kind, items[i], dt, ang, ec)
    assert n_all >= 5 and n_ok >= n_all - 1

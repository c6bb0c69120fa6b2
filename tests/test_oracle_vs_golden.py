"""Pin the numpy oracle (oracle/dspslam_oracle.py) to vectors produced by the
reference's own code (tests/golden/*.npz, made by oracle/make_goldens.py)."""
import os

import numpy as np
import pytest

from oracle import dspslam_oracle as orc
from dspslam import synth

G = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    with np.load(os.path.join(G, name)) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="module")
def dec_fit():
    return orc.OracleDecoder(synth.load_wfit().state_dict())


@pytest.fixture(scope="module")
def dec_rand():
    return orc.OracleDecoder(synth.make_decoder(0).state_dict())


@pytest.mark.parametrize("which", ["wrand", "wfit"])
def test_queries(which, dec_fit, dec_rand):
    g = load("queries_%s.npz" % which)
    dec = dec_fit if which == "wfit" else dec_rand
    sdf = dec.sdf(g["code"], g["xyz"])
    s2, jac = dec.sdf_jacobian(g["code"], g["xyz"])
    assert np.abs(sdf - g["sdf"]).max() < 1e-5      # FP32 validation tolerance (north_star)
    assert np.abs(s2 - g["sdf_j"]).max() < 1e-5
    assert np.abs(jac - g["jac"]).max() < 1e-5


def test_jacobian_finite_difference(dec_fit):
    d64 = orc.OracleDecoder(synth.load_wfit().state_dict(), dtype=np.float64)
    rng = np.random.default_rng(0)
    code = 0.2 * rng.normal(size=64)
    x = rng.uniform(-0.5, 0.5, size=(5, 3))
    _, jac = d64.sdf_jacobian(code, x)
    eps = 1e-6
    for c in range(67):
        cp, xp = code.copy(), x.copy()
        cm, xm = code.copy(), x.copy()
        if c < 64:
            cp[c] += eps
            cm[c] -= eps
        else:
            xp[:, c - 64] += eps
            xm[:, c - 64] -= eps
        fd = (d64.sdf(cp, xp) - d64.sdf(cm, xm)) / (2 * eps)
        assert np.abs(fd - jac[:, c]).max() < 1e-5


def test_liealg_huber_rot():
    g = load("liealg.npz")
    for x, s7, s6 in zip(g["x"], g["exp_sim3"], g["exp_se3"]):
        assert np.abs(orc.exp_sim3(x) - s7).max() < 2e-6
        assert np.abs(orc.exp_se3(x[:6]) - s6).max() < 2e-6
    assert np.allclose(orc.exp_sim3(np.zeros(7, np.float32)), np.eye(4))
    for b in (0.025, 0.2, 0.05):
        rr, loss, w = orc.get_robust_res(g["hub_in"].copy(), b)
        assert np.abs(rr - g["hub_r_%g" % b]).max() < 1e-7
        assert abs(loss - g["hub_loss_%g" % b]) < 1e-8
        assert np.abs(w - g["hub_w_%g" % b]).max() < 1e-6
    for T, J, r in zip(g["rot_T_oc"], g["rot_J"], g["rot_res"]):
        Jo, ro = orc.compute_rotation_loss_sim3(T)
        assert np.abs(Jo - J).max() < 2e-5 and abs(ro - r) < 2e-6


def test_exp_sim3_matches_expm_for_positive_sigma():
    from scipy.linalg import expm
    x = np.array([0.3, -0.1, 0.2, 0.2, -0.3, 0.1, 0.15])
    A = np.zeros((4, 4))
    A[:3, :3] = orc._hat(x[3:6], np.float64) + x[6] * np.eye(3)
    A[:3, 3] = x[:3]
    assert np.abs(orc.exp_sim3(x, np.float64) - expm(A)).max() < 1e-12
    # documented quirk: for sigma<0 (theta>1e-8) the reference uses c=0 -> differs from expm
    x[6] = -0.15
    A[:3, :3] = orc._hat(x[3:6], np.float64) + x[6] * np.eye(3)
    assert np.abs(orc.exp_sim3(x, np.float64) - expm(A)).max() > 1e-3


def test_voxel_grid_and_grid_sdf(dec_fit):
    g = load("voxel_grid.npz")
    from reconstruct.utils import create_voxel_grid
    for k, ref in g.items():
        G_ = int(k[1:])
        assert np.array_equal(orc.create_voxel_grid(G_), ref), k
        assert np.array_equal(create_voxel_grid(G_), ref), k
    lat = orc.create_voxel_grid(8, "lattice")
    assert np.allclose(lat[1], [-1, -1, -1 + 2 / 7]) and np.allclose(lat[8], [-1, -1 + 2 / 7, -1])
    gs = load("grid_sdf.npz")
    sdf = orc.grid_sdf(dec_fit, gs["code"], int(gs["G"]))
    assert np.abs(sdf - gs["sdf"]).max() < 1e-5


# The reference's free-running GN trajectory is chaotic at the 1e-2 level: its own
# result changes by up to 0.3 deg / 18 % code-L2 when only the torch thread count
# changes (measured, DESIGN.md "noise floor"), because the strict row-selection
# thresholds and the k4=1e7 rotation prior amplify FP32 rounding.  Parity is
# therefore pinned per step ("teacher forcing": every iteration restarts from the
# reference's recorded state) at tight tolerances, and the free run is only
# required to land near the reference.
def _one_step_cfg(cfg_fn):
    cfg = cfg_fn()
    cfg.optimizer.joint_optim.num_iterations = 1
    return cfg


def _teacher_forced(g, dec, cfg_fn, steps):
    opt = orc.OracleOptimizer(dec, _one_step_cfg(cfg_fn))
    n_it = g["H"].shape[0]
    for e in steps:
        T_co = np.linalg.inv(g["T_oc"][e].astype(np.float64)).astype(np.float32)
        trace = []
        out = opt.reconstruct_object(T_co, g["in_pts"], g["in_rays"], g["in_depth"], g["code_it"][e],
                                     trace=trace)
        H, b = trace[0]["H"], trace[0]["b"]
        assert abs(trace[0]["K"] - g["K"][e]) <= 2, e
        assert np.linalg.norm(H - g["H"][e]) / np.linalg.norm(g["H"][e]) < 3e-3, e
        assert np.linalg.norm(b - g["b"][e]) / np.linalg.norm(g["b"][e]) < 1e-2, e
        if e + 1 < n_it:
            T_next, code_next = np.linalg.inv(g["T_oc"][e + 1].astype(np.float64)), g["code_it"][e + 1]
        else:
            T_next, code_next = g["t_cam_obj"], g["code"]
        dt, ang, _ = synth.pose_errors(out["t_cam_obj"], T_next)
        assert dt < 1e-3 and ang < 0.02, (e, dt, ang)          # << 1 cm / 0.1 deg per step
        assert np.linalg.norm(out["code"] - code_next) < 1e-2 * max(np.linalg.norm(code_next), 0.1), e


def _free_run(g, out):
    dt, ang, _ = synth.pose_errors(out["t_cam_obj"], g["t_cam_obj"])
    assert dt < 0.01 and ang < 1.0   # free run: reference self-noise is up to 0.3 deg
    assert np.linalg.norm(out["code"] - g["code"]) < 0.05
    assert abs(out["loss"] - g["loss"]) < 2e-2 * max(1.0, abs(g["loss"]))


@pytest.mark.parametrize("k", [0, 1, 2, 3])
def test_c1_reconstruct_object(k, dec_fit):
    g = load("c1_obj%d.npz" % k)
    _teacher_forced(g, dec_fit, synth.kitti_config, range(10))
    if k < 2:
        opt = orc.OracleOptimizer(dec_fit, synth.kitti_config())
        out = opt.reconstruct_object(g["in_t_cam_obj"], g["in_pts"], g["in_rays"], g["in_depth"])
        assert out["is_good"] and bool(g["is_good"])
        _free_run(g, out)


@pytest.mark.parametrize("k", [0, 1])
def test_c3_mono(k, dec_fit):
    g = load("c3_obj%d.npz" % k)
    _teacher_forced(g, dec_fit, synth.redwood_config, [0, 1, 5, 10, 19])


@pytest.mark.parametrize("k", [0, 1])
def test_c3_freiburg(k, dec_fit):
    """configs/config_freiburg_001.json (k1 = 1, k3 = 0.5, scale damping 100), the reference's own 50 samples."""
    g = load("c3f_obj%d.npz" % k)
    _teacher_forced(g, dec_fit, lambda: synth.freiburg_config(iters=10, num_depth_samples=50), [0, 1, 4, 9])


@pytest.mark.parametrize("k", [0, 1])
def test_pose_only(k, dec_fit):
    g = load("pose_only_obj%d.npz" % k)
    opt = orc.OracleOptimizer(dec_fit, synth.kitti_config())
    trace = []
    T = opt.estimate_pose_cam_obj(g["in_t_co_se3"].copy(), float(g["in_scale"]), g["in_pts"], g["in_code"],
                                  trace=trace)
    for e in range(g["H"].shape[0]):
        assert np.linalg.norm(trace[e]["H"] - g["H"][e]) / np.linalg.norm(g["H"][e]) < 2e-3
        assert np.linalg.norm(trace[e]["b"] - g["b"][e]) / np.linalg.norm(g["b"][0]) < 5e-3
    assert np.abs(T - g["t_cam_obj"]).max() < 1e-3


def test_wrand_failure_path(dec_rand):
    g = load("wrand_fail.npz")
    opt = orc.OracleOptimizer(dec_rand, synth.kitti_config())
    out = opt.reconstruct_object(g["in_t_cam_obj"], g["in_pts"], g["in_rays"], g["in_depth"])
    assert not bool(g["is_good"]) and not out["is_good"] and out["t_cam_obj"] is None


def test_objectpoint(dec_fit):
    g = load("objectpoint.npz")
    opt = orc.OracleOptimizer(dec_fit, synth.kitti_config())
    assert abs(opt.compute_sdf_loss_objectpoint_zhjd(g["pts"], g["code"]) - g["mean_sdf"]) < 1e-6

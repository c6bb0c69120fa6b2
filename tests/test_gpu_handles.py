"""Handle hygiene on the GPU box: several library handles with different decoders on one device, calls of one handle
arriving on different CUDA streams, and a code_len = 32 decoder (the reference's mono caller handles both 64 and 32:
src/LocalMapping_util.cc:416-426) on both arithmetic paths."""
import threading
import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _params(cfg, iters):
    j = cfg.optimizer.joint_optim
    return dict(k1=j.k1, k2=j.k2, k3=j.k3, k4=j.k4, b1=j.b1, b2=j.b2, lr=j.learning_rate, s_damp=j.scale_damping,
                cut_off=cfg.optimizer.cut_off_threshold, num_iterations=iters,
                num_depth_samples=cfg.optimizer.num_depth_samples)


@pytest.mark.parametrize("mode", ["fp32", "tc"])
def test_two_handles_with_different_decoders_interleaved(mode):
    """The per-layer stage table of the tcgen05 kernels used to be one process-global __constant__ array: the second
    handle overwrote the first one's.  It now travels with the handle; interleaved calls equal isolated calls."""
    from dspslam import synth
    from dspslam.engine import Engine
    fit, rnd = synth.load_wfit(), synth.make_decoder(0)
    rng = np.random.default_rng(5)
    code = (0.2 * rng.normal(size=64)).astype(np.float32)
    x = rng.uniform(-0.7, 0.7, size=(3000, 3)).astype(np.float32)
    alone = []
    for dec in (fit, rnd):
        e = Engine(dec)
        s, j = e.sdf_jac(code, x, mode=mode)
        alone.append((e.sdf(code, x, mode=mode).cpu().numpy(), s.cpu().numpy(), j.cpu().numpy()))
        e.close()
    ea, eb = Engine(fit), Engine(rnd)          # b created after a: a's table must survive
    for _ in range(2):
        for e, ref in ((ea, alone[0]), (eb, alone[1])):
            s, j = e.sdf_jac(code, x, mode=mode)
            assert np.array_equal(e.sdf(code, x, mode=mode).cpu().numpy(), ref[0])
            assert np.array_equal(s.cpu().numpy(), ref[1]) and np.array_equal(j.cpu().numpy(), ref[2])
    assert not np.allclose(alone[0][0], alone[1][0])


def _decoder32(seed=3):
    import torch
    from deep_sdf.deep_sdf_decoder import Decoder
    from dspslam import synth
    torch.manual_seed(seed)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        dec = Decoder(32, **synth.NETWORK_SPECS).eval()
    # a default-initialised decoder is almost constant (SDF 0.029 +- 0.001): stretch the output layer and centre it so
    # that the SDF crosses zero inside the unit ball and a GN step has render rows
    from oracle import dspslam_oracle as orc
    with torch.no_grad():
        dec.lin8.weight.mul_(40.0)
        x = np.random.default_rng(1).uniform(-0.6, 0.6, size=(4000, 3)).astype(np.float32)
        pre = np.arctanh(orc.OracleDecoder(dec.state_dict()).sdf(np.zeros(32, np.float32), x))
        dec.lin8.bias.sub_(float(np.median(pre)))
    return dec


@pytest.mark.parametrize("mode", ["fp32", "tc"])
def test_code_len_32_queries_vs_oracle(mode):
    from dspslam.engine import Engine
    from oracle import dspslam_oracle as orc
    dec = _decoder32()
    eng = Engine(dec)
    assert eng.code_len == 32
    od = orc.OracleDecoder(dec.state_dict())
    rng = np.random.default_rng(11)
    code = (0.3 * rng.normal(size=32)).astype(np.float32)
    x = rng.uniform(-0.8, 0.8, size=(5000, 3)).astype(np.float32)
    so, jo = od.sdf_jacobian(code, x)
    tol = 1e-5 if mode == "fp32" else 1e-3
    assert np.abs(eng.sdf(code, x, mode=mode).cpu().numpy() - so).max() < tol
    s, j = eng.sdf_jac(code, x, mode=mode)
    assert j.shape == (5000, 35)
    assert np.abs(s.cpu().numpy() - so).max() < tol
    err = np.abs(j.cpu().numpy() - jo)
    if mode == "fp32":
        assert (err.max(axis=1) >= 1e-5).sum() <= 2 and err.max() < 5e-2
    else:
        assert np.sqrt((err ** 2).mean()) < 6e-4
    g = eng.grid_sdf(code, 12, coord_mode="lattice", mode=mode).cpu().numpy().reshape(12, 12, 12)
    assert np.abs(g - orc.grid_sdf(od, code, 12, "lattice")).max() < tol


def test_code_len_32_exact_mode_vs_oracle():
    """The hi/lo kernel with a 32-long code: 35 Jacobian columns (the last step's N is 48), layer 3 emits 477 units."""
    from dspslam.engine import Engine
    from oracle import dspslam_oracle as orc
    dec = _decoder32()
    eng = Engine(dec)
    od = orc.OracleDecoder(dec.state_dict())
    rng = np.random.default_rng(12)
    code = (0.3 * rng.normal(size=32)).astype(np.float32)
    x = rng.uniform(-0.8, 0.8, size=(3001, 3)).astype(np.float32)
    so, jo = od.sdf_jacobian(code, x)
    s, j = eng.sdf_jac(code, x, mode="tc_exact")
    assert j.shape == (3001, 35)
    assert np.abs(s.cpu().numpy() - so).max() < 2e-5
    err = np.abs(j.cpu().numpy() - jo)
    assert (err.max(axis=1) > 1e-4).sum() <= 12 and np.median(err.max(axis=1)) < 2e-5


def test_code_len_32_gn_step_vs_oracle():
    """One joint GN step (39 x 39 normal equations) in FP32 mode against the oracle; the tensor-core mode against
    the FP32 mode."""
    from dspslam import synth
    from dspslam.engine import Engine
    from oracle import dspslam_oracle as orc
    from reconstruct.utils import ForceKeyErrorDict
    dec = _decoder32()
    eng = Engine(dec)
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.k4 = 0.0
    cfg.optimizer.code_len = 32
    cfg.optimizer.joint_optim.num_iterations = 1
    rng = np.random.default_rng(2)
    # object at 6 m in front of the camera, unit scale; surface points and rays through the unit ball
    T = np.eye(4, dtype=np.float32)
    T[:3, :3] = np.diag([1, -1, -1]).astype(np.float32)
    T[:3, 3] = [0.2, 0.1, 6.0]
    po = rng.uniform(-0.5, 0.5, size=(300, 3)).astype(np.float32)
    pts = (po @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    rays = np.concatenate([pts / pts[:, 2:3], (rng.uniform(-0.3, 0.3, size=(120, 3)) + [0, 0, 1]).astype(np.float32)])
    rays = (rays / rays[:, 2:3]).astype(np.float32)
    depth = pts[:, 2].copy()
    code = (0.1 * rng.normal(size=32)).astype(np.float32)
    ref = orc.OracleOptimizer(orc.OracleDecoder(dec.state_dict()), cfg).reconstruct_object(T, pts, rays, depth, code=code)
    res = eng.gn_joint([(T, pts, rays, depth, code)], _params(cfg, 1), mode="fp32")
    assert bool(ref["is_good"]) == (res.status[0] == 0)
    if ref["is_good"]:
        dt, ang, _ = synth.pose_errors(res.t_cam_obj[0], ref["t_cam_obj"])
        assert res.code.shape == (1, 32)
        assert dt < 1e-3 and ang < 0.05 and np.linalg.norm(res.code[0] - ref["code"]) < 1e-2 * max(1.0, np.linalg.norm(ref["code"]))
        for m, tol in (("tc", 0.5), ("tc_exact", 0.05)):
            tc = eng.gn_joint([(T, pts, rays, depth, code)], _params(cfg, 1), mode=m)
            dt, ang, _ = synth.pose_errors(tc.t_cam_obj[0], res.t_cam_obj[0])
            assert tc.status[0] == 0 and dt < 1e-2 and ang < tol, (m, dt, ang)


@pytest.mark.parametrize("mode", ["fp32", "tc"])
def test_one_handle_two_streams_two_threads(mode):
    """Two host threads, each on its own CUDA stream, share one handle (scratch, latent staging, extras record):
    the library orders the calls on the device with an event, so both get the results of a serial run."""
    import torch
    from dspslam import synth
    from dspslam.engine import Engine
    eng = Engine(synth.load_wfit())
    prm = _params(synth.kitti_config(), 2)
    dets = [[synth.make_kitti_detection(10 * t + k, M=150 + 40 * k, n_bg=60) for k in range(3)] for t in range(2)]
    rng = np.random.default_rng(0)
    codes = [(0.2 * rng.normal(size=64)).astype(np.float32) for _ in range(2)]
    xs = [rng.uniform(-0.7, 0.7, size=(4000 + 1000 * t, 3)).astype(np.float32) for t in range(2)]
    serial = []
    for t in range(2):
        r = eng.gn_joint(dets[t], prm, mode=mode)
        serial.append((r.t_cam_obj.copy(), r.code.copy(), eng.sdf(codes[t], xs[t], mode=mode).cpu().numpy()))
    out, errs = [None, None], []

    def work(t):
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for _ in range(6):
                    r = eng.gn_joint(dets[t], prm, mode=mode)
                    s = eng.sdf(codes[t], xs[t], mode=mode)
                    st.synchronize()
                    out[t] = (r.t_cam_obj, r.code, s.cpu().numpy())
                    for a, b in zip(out[t], serial[t]):
                        assert np.array_equal(a, b)
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)

    th = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errs, errs


@pytest.mark.parametrize("mode", ["tc", "tc_fast"])
def test_early_jacobian_tiles_on_second_stream_bit_identical(mode):
    """Experiment knob DSPSLAM_JAC_EARLY=1: the Jacobian tiles made of surface rows only run on the handle's second
    stream beside the ray pipeline of the same iteration (api.cu: gn_run; off by default, measured neutral).  Same
    tiles, same partial slots, same summation order: the result is bit-identical to the single-stream schedule, run
    after run."""
    import os
    from dspslam import synth
    from dspslam.engine import Engine
    dets = [synth.make_kitti_detection(40 + k, M=700 + 130 * k, n_bg=80) for k in range(5)]   # 5+6+7+8+9 = 35 early tiles
    prm = _params(synth.kitti_config(), 4)
    old = os.environ.get("DSPSLAM_JAC_EARLY")
    try:
        os.environ["DSPSLAM_JAC_EARLY"] = "0"
        e0 = Engine(synth.load_wfit())
        os.environ["DSPSLAM_JAC_EARLY"] = "1"
        e1 = Engine(synth.load_wfit())
        os.environ.pop("DSPSLAM_JAC_EARLY")
    finally:
        if old is not None:
            os.environ["DSPSLAM_JAC_EARLY"] = old
    a = e0.gn_joint(dets, prm, mode=mode, debug=True)
    for _ in range(3):
        b = e1.gn_joint(dets, prm, mode=mode, debug=True)
        assert np.array_equal(a.t_cam_obj, b.t_cam_obj) and np.array_equal(a.code, b.code)
        assert np.array_equal(a.dbg, b.dbg) and np.array_equal(a.status, b.status)
    assert (a.status == 0).all()


def test_query_reports_the_knobs_and_a_bad_mask_is_refused(monkeypatch):
    """dspslam_query gives back what the environment variables of dspslam_create selected (include/dspslam.h); a mask
    that does not parse fails the creation instead of silently selecting the tc_fast Jacobian arithmetic."""
    from dspslam import _native as N
    from dspslam import synth
    from dspslam.engine import Engine
    monkeypatch.delenv("DSPSLAM_WSPLIT_MASK", raising=False)
    monkeypatch.delenv("DSPSLAM_JAC_EARLY", raising=False)
    e = Engine(synth.load_wfit())
    assert e.lib.dspslam_query(e._h, N.Q_WSPLIT_MASK) == 0xF0
    assert e.lib.dspslam_query(e._h, N.Q_JAC_EARLY) == 0
    assert e.lib.dspslam_query(e._h, N.Q_DEVICE) == e.device.index
    assert e.lib.dspslam_query(e._h, 99) == -1 and e.lib.dspslam_query(None, 0) == -1
    monkeypatch.setenv("DSPSLAM_WSPLIT_MASK", "0x0E")
    monkeypatch.setenv("DSPSLAM_JAC_EARLY", "1")
    e2 = Engine(synth.load_wfit())
    assert e2.lib.dspslam_query(e2._h, N.Q_WSPLIT_MASK) == 0x0E and e2.lib.dspslam_query(e2._h, N.Q_JAC_EARLY) == 1
    monkeypatch.delenv("DSPSLAM_JAC_EARLY")
    for bad in ("oxF0", "", "300", "-1", "12junk"):
        monkeypatch.setenv("DSPSLAM_WSPLIT_MASK", bad)
        with pytest.raises(N.NativeError, match="DSPSLAM_WSPLIT_MASK"):
            Engine(synth.load_wfit())

"""Detection -> optimiser-input builder: reconstruct/kitti_sequence.py:112-216 and
mono_sequence.py:80-114.

CPU tests pin the numpy oracle to golden vectors produced by the reference's own classes
(oracle/make_frame_goldens.py); GPU tests require the CUDA builder (csrc/frame.cu, through the C ABI) to be
BIT-IDENTICAL to those vectors - this is index / byte work - and check the frame -> Gauss-Newton hand-over on
the device against the host path."""
import ast
import os

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
G = os.path.join(ROOT, "tests", "golden")
KITTI = [0, 1, 2]


def load(name):
    with np.load(os.path.join(G, name)) as z:
        return {k: z[k] for k in z.files}


def frame(seed):
    from dspslam.frame_synth import synth_kitti_frame
    g = load("frame_kitti_%d.npz" % seed)
    return g, synth_kitti_frame(seed, **ast.literal_eval(str(g["kw"])))


def check_instances(inst, g, get):
    assert len(inst) == int(g["n_inst"])
    for i, ins in enumerate(inst):
        assert np.array_equal(get(ins, "T_cam_obj"), g["i%d_T_cam_obj" % i])
        assert np.array_equal(get(ins, "surface_points"), g["i%d_pts" % i])
        assert bool(get(ins, "is_front")) == bool(g["i%d_is_front" % i])
        rays = get(ins, "rays")
        assert (rays is not None) == bool(g["i%d_has_rays" % i])
        if rays is not None:
            assert np.array_equal(rays, g["i%d_rays" % i])
            assert np.array_equal(get(ins, "depth"), g["i%d_depth" % i])


# ---- CPU: oracle against the reference's vectors ---------------------------------------------------------------
@pytest.mark.parametrize("seed", KITTI)
def test_oracle_kitti_frame_vs_reference_golden(seed):
    from dspslam.frame_synth import IMG_H, IMG_W, K, T_CAM_VELO
    from oracle import frame_oracle as fo
    g, (scan, boxes, masks, bbs) = frame(seed)
    inst = fo.kitti_frame(scan, boxes, masks, bbs, K, np.linalg.inv(K).astype(np.float32), T_CAM_VELO, IMG_W, IMG_H)
    check_instances(inst, g, lambda ins, k: ins[k])
    for i, ins in enumerate(inst):
        assert ins["mask_id"] == int(g["i%d_mask_id" % i])


@pytest.mark.parametrize("seed", [0, 1])
def test_oracle_mono_background_rays_vs_reference_golden(seed):
    from dspslam.frame_synth import IMG_H, IMG_W, synth_kitti_frame
    from oracle import frame_oracle as fo
    g = load("frame_mono_%d.npz" % seed)
    _, _, masks, bbs = synth_kitti_frame(int(g["seed"]), n_cars=3, n_ground=10)
    n, rays = fo.mono_background_rays(masks, bbs, g["K"], np.linalg.inv(g["K"]), float(g["k1"]), float(g["k2"]),
                                      IMG_W, IMG_H)
    assert n == int(g["mask_id"]) and np.array_equal(rays, g["background_rays"])


def test_linspace_restatement_matches_numpy():
    from oracle import frame_oracle as fo
    rng = np.random.default_rng(0)
    for _ in range(300):
        a, n = int(rng.integers(0, 50)), int(rng.integers(1, 400))
        b = a + int(rng.integers(0, 5000))
        assert np.array_equal(fo.linspace_int(a, b, n), np.linspace(a, b, n).astype(np.int32))


def test_undistort_restatement_known_answers():
    """No distortion is the identity; the fixed point satisfies the forward radial model to ~1e-3 pixel."""
    from oracle import frame_oracle as fo
    K = np.array([[535.4, 0, 320.1], [0, 539.2, 247.6], [0, 0, 1]])
    pix = np.array([[10, 20], [320, 240], [600, 400]], np.float32)
    assert np.allclose(fo.undistort_points(pix, K, 0.0, 0.0), pix, atol=1e-4)
    und = fo.undistort_points(pix, K, -0.1, 0.02).astype(np.float64)
    x, y = (und[:, 0] - K[0, 2]) / K[0, 0], (und[:, 1] - K[1, 2]) / K[1, 1]
    r2 = x * x + y * y
    f = 1 + (-0.1) * r2 + 0.02 * r2 * r2
    assert np.allclose(np.stack([x * f * K[0, 0] + K[0, 2], y * f * K[1, 1] + K[1, 2]], 1), pix, atol=5e-3)


# ---- GPU: the CUDA builder, bit-exact -----------------------------------------------------------------------------
@pytest.fixture(scope="module")
def eng():
    from dspslam import synth
    from dspslam.engine import Engine
    return Engine(synth.load_wfit())


def _builder(eng, **kw):
    from dspslam.frame import FrameBuilder
    from dspslam.frame_synth import IMG_H, IMG_W, K, T_CAM_VELO
    return FrameBuilder(eng, K, np.linalg.inv(K).astype(np.float32), T_CAM_VELO, IMG_W, IMG_H, **kw)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", KITTI)
def test_gpu_kitti_frame_bit_exact_vs_reference_golden(seed, eng):
    g, (scan, boxes, masks, bbs) = frame(seed)
    b = _builder(eng)
    fb = b.build_kitti(scan, boxes, masks, bbs)
    inst = b.instances(fb, masks, bbs)
    check_instances(inst, g, lambda ins, k: ins[k])
    for i, ins in enumerate(inst):
        assert int(fb.info[i, 3]) == int(g["i%d_mask_id" % i])
        assert ins.num_surface_points == g["i%d_pts" % i].shape[0]
    # packed layout invariants of the dspslam_gn_joint hand-over
    assert fb.pts_off[0] == 0 and fb.ray_off[0] == 0
    assert np.all(np.diff(fb.pts_off) == fb.info[:, 1])
    assert np.all(np.diff(fb.ray_off) == np.where(fb.info[:, 4] > 0, fb.info[:, 1] + fb.info[:, 5], 0))


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(num_lidar_max=1), dict(num_lidar_max=7, max_background=3, downsample_ratio=1.0),
                                 dict(num_lidar_max=100000, max_background=1024, downsample_ratio=2.0),
                                 dict(num_lidar_max=64, min_mask_area=10 ** 7)])
def test_gpu_kitti_frame_vs_oracle_edge_configs(cfg, eng):
    """Sub-sampling caps of 1 / tiny / never reached, every lattice pixel (stride 1), masks all too small."""
    from dspslam.frame_synth import IMG_H, IMG_W, K, T_CAM_VELO, synth_kitti_frame
    from oracle import frame_oracle as fo
    scan, boxes, masks, bbs = synth_kitti_frame(5, n_cars=4, n_ground=30000, pts_per_car=900)
    ref = fo.kitti_frame(scan, boxes, masks, bbs, K, np.linalg.inv(K).astype(np.float32), T_CAM_VELO, IMG_W, IMG_H,
                         num_lidar_max=cfg.get("num_lidar_max", 250), min_mask_area=cfg.get("min_mask_area", 1000),
                         downsample_ratio=cfg.get("downsample_ratio", 4.0), max_bg=cfg.get("max_background", 200))
    b = _builder(eng, **cfg)
    inst = b.instances(b.build_kitti(scan, boxes, masks, bbs))
    assert len(inst) == len(ref)
    for a, r in zip(inst, ref):
        assert np.array_equal(a.surface_points, r["surface_points"])
        assert (a.rays is None) == (r["rays"] is None)
        if a.rays is not None:
            assert np.array_equal(a.rays, r["rays"]) and np.array_equal(a.depth, r["depth"])


@pytest.mark.gpu
def test_gpu_frame_empty_inputs(eng):
    """No LiDAR points, no masks: detections keep zero points / no rays (kitti_sequence.py:173-174)."""
    from dspslam.frame_synth import synth_kitti_frame
    scan, boxes, masks, bbs = synth_kitti_frame(6, n_cars=2, n_ground=2000)
    b = _builder(eng)
    fb = b.build_kitti(scan, boxes, masks[:0], bbs[:0])
    assert np.all(fb.info[:, 3] == -1) and fb.total_rays == 0 and fb.total_pts > 0
    fb = b.build_kitti(scan[:0], boxes, masks, bbs)
    assert fb.total_pts == 0 and fb.total_rays == 0
    # more than 32 detections: the crop kernels run in groups of 32 boxes
    many = np.concatenate([boxes] * 13)[:37]
    fb = b.build_kitti(scan, many, masks, bbs)
    inst = b.instances(fb)
    one = b.instances(b.build_kitti(scan, boxes, masks, bbs))
    assert len(inst) == 37 and fb.info[:, 1].sum() == fb.total_pts
    for i in inst:                           # every copy equals the instance of the box it duplicates
        src = [o for o in one if np.array_equal(o.T_cam_obj, i.T_cam_obj)]
        assert len(src) == 1 and np.array_equal(i.surface_points, src[0].surface_points)
        assert (i.rays is None) == (src[0].rays is None)
        if i.rays is not None:
            assert np.array_equal(i.rays, src[0].rays)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [0, 1])
def test_gpu_mono_background_rays_bit_exact_vs_reference_golden(seed, eng):
    from dspslam.frame import FrameBuilder
    from dspslam.frame_synth import IMG_H, IMG_W, synth_kitti_frame
    g = load("frame_mono_%d.npz" % seed)
    _, _, masks, bbs = synth_kitti_frame(int(g["seed"]), n_cars=3, n_ground=10)
    b = FrameBuilder(eng, g["K"], None, None, IMG_W, IMG_H, k1=float(g["k1"]), k2=float(g["k2"]))
    mid, rays = b.background_rays_mono(masks, bbs)
    assert mid == int(g["mask_id"]) and np.array_equal(rays, g["background_rays"])


@pytest.mark.gpu
def test_gpu_frame_to_gauss_newton_on_device(eng):
    """The device-resident hand-over: reconstruct_frame(FrameBatch) == reconstruct_objects(host copies of the same
    detections), and the reference-shaped sequence mirror returns the same instances."""
    import types
    from dspslam import synth
    from dspslam.frame_synth import K, T_CAM_VELO
    from reconstruct.kitti_sequence import FrameWithLiDAR
    from reconstruct.optimizer import Optimizer
    g, (scan, boxes, masks, bbs) = frame(2)
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = 2
    opt = Optimizer(synth.load_wfit(), cfg, mode="fp32")
    b = _builder(opt.engine)
    fb = b.build_kitti(scan, boxes, masks, bbs)
    inst = b.instances(fb)
    outs = opt.reconstruct_frame(fb)
    assert len(outs) == fb.B
    dets = [(i.T_cam_obj, i.surface_points, i.rays, i.depth) for i in inst if i.rays is not None]
    refs = iter(opt.reconstruct_objects(dets))
    for i, o in zip(inst, outs):
        if i.rays is None:
            assert not o.is_good
            continue
        r = next(refs)
        assert o.is_good == r.is_good
        if o.is_good:
            assert np.allclose(o.t_cam_obj, r.t_cam_obj, atol=1e-5) and np.allclose(o.code, r.code, atol=1e-5)
    # the reference-shaped mirror (arrays instead of files)
    seq = types.SimpleNamespace(configs=synth.kitti_config(), K_cam=K, invK_cam=np.linalg.inv(K).astype(np.float32),
                                T_cam_velo=T_CAM_VELO, engine=opt.engine)
    fr = FrameWithLiDAR(seq, 0, velo_pts=scan, detections_3d=boxes, det_2d={"pred_masks": masks, "pred_boxes": bbs})
    fr.get_detections()
    check_instances(fr.instances, g, lambda ins, k: ins[k])


# ---- CPU: host-side logic of the builder and the boundary structs --------------------------------------------------
def test_box_frames_host_bookkeeping_matches_oracle():
    from dspslam.frame import box_frames
    from dspslam.frame_synth import T_CAM_VELO, synth_kitti_frame
    from oracle import frame_oracle as fo
    _, boxes, _, _ = synth_kitti_frame(4, n_cars=6, n_ground=100)
    order, cb, T_co, sizes = box_frames(boxes, T_CAM_VELO)
    oorder, ob = fo.box_frames(boxes, T_CAM_VELO)
    assert np.array_equal(order, oorder)
    for i, o in enumerate(ob):
        assert np.array_equal(np.array(cb[i].T_obj_velo[:], np.float32).reshape(3, 4), o["T_obj_velo"][:3])
        assert np.array_equal(np.array(cb[i].lo[:], np.float32), o["lo"])
        assert np.array_equal(np.array(cb[i].hi[:], np.float32), o["hi"])
        assert np.array_equal(np.array(cb[i].half[:], np.float32), o["half"])
        assert np.array_equal(T_co[i], o["T_cam_obj"]) and np.array_equal(sizes[i], o["scale"])
        assert bool(cb[i].is_front) == bool(o["T_cam_obj"][2, 3] > 0)


def test_frame_structs_match_the_c_header(tmp_path):
    """ctypes mirrors of dspslam_frame_params / dspslam_frame_box have the size and field offsets gcc gives the
    declarations in include/dspslam.h."""
    import ctypes
    import subprocess
    from dspslam import _native as N
    src = tmp_path / "s.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "dspslam.h"\nint main(void){'
                   'printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(dspslam_frame_params), offsetof(dspslam_frame_params, K),'
                   'offsetof(dspslam_frame_params, T_cam_velo), offsetof(dspslam_frame_params, k1),'
                   'offsetof(dspslam_frame_params, invK64), sizeof(dspslam_frame_box), offsetof(dspslam_frame_box, is_front));'
                   'printf("%zu %zu\\n", sizeof(dspslam_joint_params), sizeof(dspslam_profile));return 0;}')
    exe = tmp_path / "s"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    a, b = subprocess.check_output([str(exe)]).decode().split("\n")[:2]
    got = [int(v) for v in a.split()]
    P, B = N.FrameParams, N.FrameBox
    assert got == [ctypes.sizeof(P), P.K.offset, P.T_cam_velo.offset, P.k1.offset, P.invK64.offset,
                   ctypes.sizeof(B), B.is_front.offset]
    assert [int(v) for v in b.split()] == [ctypes.sizeof(N.JointParams), ctypes.sizeof(N.Profile)]


def test_kitti_sequence_calibration(tmp_path):
    """KITIISequence.load_calib (kitti_sequence.py:243-257): K = P2[:, :3], T_cam_velo = [I | (P2[0,3]/fx, 0, 0)] Tr."""
    import types
    from reconstruct.kitti_sequence import KITIISequence
    (tmp_path / "calib.txt").write_text(
        "P0: 7.18856e+02 0 6.071928e+02 0 0 7.18856e+02 1.852157e+02 0 0 0 1 0\n"
        "P2: 7.18856e+02 0 6.071928e+02 4.538225e+01 0 7.18856e+02 1.852157e+02 -1.130887e-01 0 0 1 3.779761e-03\n"
        "Tr: 4.276802e-04 -9.999672e-01 -8.084491e-03 -1.198459e-02 -7.210626e-03 8.081198e-03 -9.999413e-01 "
        "-5.403984e-02 9.999738e-01 4.859485e-04 -7.206933e-03 -2.921968e-01\n")
    seq = KITIISequence.__new__(KITIISequence)
    seq.calib_file = str(tmp_path / "calib.txt")
    seq.load_calib()
    assert seq.K_cam.dtype == np.float32 and seq.T_cam_velo.dtype == np.float32
    assert np.allclose(seq.K_cam, [[718.856, 0, 607.1928], [0, 718.856, 185.2157], [0, 0, 1]])
    assert np.allclose(seq.K_cam.astype(np.float64) @ seq.invK_cam.astype(np.float64), np.eye(3), atol=1e-4)
    Tr = np.array([4.276802e-04, -9.999672e-01, -8.084491e-03, -1.198459e-02, -7.210626e-03, 8.081198e-03, -9.999413e-01,
                   -5.403984e-02, 9.999738e-01, 4.859485e-04, -7.206933e-03, -2.921968e-01]).reshape(3, 4)
    want = np.vstack([Tr, [0, 0, 0, 1]])
    want[0, 3] += 4.538225e+01 / 7.18856e+02
    assert np.allclose(seq.T_cam_velo, want, atol=1e-6)
    _ = types


@pytest.mark.gpu
def test_gpu_frame_pose_only_on_resident_points(eng):
    """Re-observation path (src/LocalMapping_util.cc:94-111): pose-only GN on the surface points the builder left in
    HBM equals the host-array path `estimate_poses_cam_obj` on copies of the same points."""
    from dspslam import synth
    from reconstruct.optimizer import Optimizer
    g, (scan, boxes, masks, bbs) = frame(0)
    opt = Optimizer(synth.load_wfit(), synth.kitti_config(), mode="fp32")
    b = _builder(opt.engine)
    fb = b.build_kitti(scan, boxes, masks, bbs)
    inst = b.instances(fb)
    T, sc = [], []
    for i in inst:
        s = float(np.cbrt(np.linalg.det(i.T_cam_obj[:3, :3].astype(np.float64))))
        se3 = i.T_cam_obj.copy()
        se3[:3, :3] /= np.float32(s)
        T.append(se3)
        sc.append(s)
    codes = np.zeros((fb.B, 64), np.float32)
    codes[:, :3] = 0.1
    got = opt.estimate_poses_frame(fb, np.stack(T), sc, codes)
    want = opt.estimate_poses_cam_obj([(T[k], sc[k], inst[k].surface_points, codes[k]) for k in range(fb.B)])
    assert got.shape == (fb.B, 4, 4) and np.isfinite(got).all()
    assert np.allclose(got, want, atol=1e-6)
    assert np.abs(got - np.stack(T)).max() > 1e-4            # the optimisation moved the poses


@pytest.mark.gpu
def test_gpu_reconstruct_frame_script(tmp_path):
    """The headless `reconstruct_frame.py` (reference: reconstruct_frame.py:26-89): frame -> detections -> GN ->
    meshes, end to end on the device; writes one PLY per reconstructed detection."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("rf_script", os.path.join(ROOT, "dsp-slam-rgbd_b200", "reconstruct_frame.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = tmp_path / "objs"
    objects = mod.main(["--synthetic", "2", "-o", str(out), "--voxels_dim", "32"])
    g = load("frame_kitti_2.npz")
    n_rays = sum(bool(g["i%d_has_rays" % i]) for i in range(int(g["n_inst"])))
    assert 1 <= len(objects) <= n_rays
    for k, det, obj in objects:
        raw = (out / ("%d.ply" % k)).read_bytes()
        assert raw.startswith(b"ply\n") and len(raw) > 1000
        assert obj.t_cam_obj.shape == (4, 4) and np.isfinite(obj.t_cam_obj).all()
    assert len((out / "objects.txt").read_text().strip().split("\n")) == len(objects)

"""The drop-in boundary, replayed the way the reference's C++ drives it:

  src/System.cc:93-99        sys.path, reconstruct.utils.get_configs(json), get_decoder(cfg),
                             import("reconstruct").attr("get_sequence")(strSequencePath, pyCfg)
  src/LocalMapping.cc:38-40  reconstruct.optimizer.Optimizer(pyDecoder, pyCfg),
                             MeshExtractor(pyDecoder, pyCfg.optimizer.code_len, pyCfg.voxels_dim)
  src/Tracking_util.cc:35-49 pySequence.get_frame_by_id(frame_id) -> det.surface_points / T_cam_obj / rays / depth
  src/LocalMapping_util.cc   reconstruct_object / estimate_pose_cam_obj / extract_mesh_from_code

Positional arguments only; arrays handed over as pybind11's Eigen casters make them (float32, Fortran order).
The CPU half checks construction (nothing touches the GPU before the first frame); the GPU half runs the calls
against a scratch sequence directory (synthetic scan, pre-stored labels in the reference's torch-pickle format).
"""
import importlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF_CFG = dict(  # configs/config_kitti.json of the reference, paths redirected to the scratch directory
    data_type="KITTI", detect_online=False, min_mask_area=1000, downsample_ratio=4, num_lidar_max=250,
    num_lidar_min=10, voxels_dim=32,
    optimizer=dict(code_len=64, num_depth_samples=50, cut_off_threshold=0.01,
                   joint_optim=dict(k1=1.0, k2=100.0, k3=0.25, k4=1e7, b1=0.20, b2=0.025, num_iterations=10,
                                    learning_rate=1.0, scale_damping=1.0),
                   pose_only_optim=dict(num_iterations=5, learning_rate=1.0)))


def _scratch_kitti(tmp, n_frames=1):
    import torch
    from dspslam import synth
    from dspslam.frame_synth import synth_kitti_frame
    seq = os.path.join(tmp, "07")
    for d in ("image_2", "velodyne"):
        os.makedirs(os.path.join(seq, d), exist_ok=True)
    l2, l3 = os.path.join(tmp, "labels2d"), os.path.join(tmp, "labels3d")
    os.makedirs(l2, exist_ok=True)
    os.makedirs(l3, exist_ok=True)
    from dspslam.frame_synth import K, T_CAM_VELO
    P2 = np.hstack([K.astype(np.float64), np.zeros((3, 1))])            # camera 2 = camera 0 for the synthetic rig
    with open(os.path.join(seq, "calib.txt"), "w") as fh:              # KITTI odometry calib.txt layout
        fh.write("P0: " + " ".join("%.12e" % v for v in P2.reshape(-1)) + "\n")
        fh.write("P2: " + " ".join("%.12e" % v for v in P2.reshape(-1)) + "\n")
        fh.write("Tr: " + " ".join("%.12e" % v for v in T_CAM_VELO[:3].astype(np.float64).reshape(-1)) + "\n")
    for f in range(n_frames):
        scan, boxes, masks, bbs = synth_kitti_frame(f)
        scan.astype(np.float32).tofile(os.path.join(seq, "velodyne", "%06d.bin" % f))
        open(os.path.join(seq, "image_2", "%06d.png" % f), "wb").close()     # only counted (kitti_sequence.py:225)
        torch.save(boxes, os.path.join(l3, "%06d.lbl" % f))
        torch.save({"pred_masks": masks, "pred_boxes": bbs}, os.path.join(l2, "%06d.lbl" % f))
    exp = os.path.join(tmp, "deepsdf_cars")
    synth.write_experiment_dir(synth.load_wfit(), exp)
    cfg = dict(REF_CFG, DeepSDF_DIR=exp, path_label_2d=l2, path_label_3d=l3)
    path = os.path.join(tmp, "config_kitti.json")
    with open(path, "w") as fh:
        json.dump(cfg, fh)
    return seq, path


def test_get_sequence_kitti_two_positional_arguments(tmp_path):
    """System.cc:95-99 up to the sequence object; no CUDA call may happen before the first frame."""
    seq_dir, cfg_path = _scratch_kitti(str(tmp_path))
    io_utils = importlib.import_module("reconstruct.utils")
    pyCfg = io_utils.get_configs(cfg_path)
    pyDecoder = io_utils.get_decoder(pyCfg)
    assert io_utils.get_decoder(pyCfg) is pyDecoder            # one module per experiment directory
    pySequence = importlib.import_module("reconstruct").get_sequence(seq_dir, pyCfg)
    assert type(pySequence).__name__ == "KITIISequence" and pySequence.num_frames == 1
    assert pySequence.K_cam.dtype == np.float32 and pySequence.T_cam_velo.shape == (4, 4)
    assert pySequence.detector_2d is None and pySequence.detector_3d is None
    assert pySequence._engine is None                          # resolved lazily
    assert importlib.import_module("reconstruct").get_detectors(pyCfg) == (None, None)


def test_get_sequence_mono_reads_orbslam_yaml(tmp_path):
    """mono_sequence.py:117-134: intrinsics and k1, k2 come from configs.slam_config_path (cv2.FileStorage)."""
    yaml = os.path.join(str(tmp_path), "freiburg_001.yaml")
    with open(yaml, "w") as fh:
        fh.write("%YAML:1.0\n\n# Camera Parameters\nCamera.fx: 535.4\nCamera.fy: 539.2\nCamera.cx: 320.1\n"
                 "Camera.cy: 247.6\n\nCamera.k1: 0.26383\nCamera.k2: -0.953104\nCamera.p1: -0.005358\n"
                 "Camera.fps: 30.0\nORBextractor.nFeatures: 2000\n")
    from reconstruct.utils import ForceKeyErrorDict
    import reconstruct
    from reconstruct.mono_sequence import read_camera_yaml
    cfg = ForceKeyErrorDict(data_type="Freiburg", detect_online=False, slam_config_path=yaml,
                            path_label_2d=str(tmp_path), min_mask_area=1000, downsample_ratio=4)
    seq = reconstruct.get_sequence(str(tmp_path), cfg)
    assert type(seq).__name__ == "MonoSequence" and seq.detector_2d is None
    want = np.array([[535.4, 0, 320.1], [0, 539.2, 247.6], [0, 0, 1]])
    assert np.array_equal(seq.K_cam, want) and seq.K_cam.dtype == np.float64
    assert np.allclose(seq.invK_cam @ want, np.eye(3)) and seq.k1 == 0.26383 and seq.k2 == -0.953104
    # the flat-line fallback reader gives the same numbers as cv2.FileStorage
    real_cv2 = sys.modules.get("cv2")
    sys.modules["cv2"] = None
    try:
        cam = read_camera_yaml(yaml)
    finally:
        if real_cv2 is not None:
            sys.modules["cv2"] = real_cv2
        else:
            sys.modules.pop("cv2", None)
    assert [cam["Camera.fx"], cam["Camera.cy"], cam["Camera.k2"]] == [535.4, 247.6, -0.953104]
    assert reconstruct.get_sequence(str(tmp_path), ForceKeyErrorDict(data_type="TUM")) is None   # as the reference


def test_unmirrored_reference_modules_still_resolve(tmp_path, monkeypatch):
    """Modules the drop-in does not mirror (detector2d, loss, ...) are found in a reference checkout further down
    sys.path (`src/System.cc:94` appends "./")."""
    other = os.path.join(str(tmp_path), "checkout", "reconstruct")
    os.makedirs(other)
    open(os.path.join(other, "__init__.py"), "w").close()
    with open(os.path.join(other, "detector9d.py"), "w") as fh:
        fh.write("WHO = 'reference checkout'\n")
    with open(os.path.join(other, "optimizer.py"), "w") as fh:
        fh.write("WHO = 'must stay shadowed'\n")
    monkeypatch.syspath_prepend(os.path.dirname(other))
    import reconstruct
    monkeypatch.setattr(reconstruct, "__path__", list(reconstruct.__path__))
    reconstruct._extend_path()
    assert importlib.import_module("reconstruct.detector9d").WHO == "reference checkout"
    assert hasattr(importlib.import_module("reconstruct.optimizer"), "MeshExtractor")
    sys.modules.pop("reconstruct.detector9d", None)


@pytest.mark.gpu
def test_replay_cpp_call_sequence_on_scratch_sequence(tmp_path):
    seq_dir, cfg_path = _scratch_kitti(str(tmp_path))
    F = np.asfortranarray
    # ---- src/System.cc:93-99 ----
    io_utils = importlib.import_module("reconstruct.utils")
    pyCfg = io_utils.get_configs(cfg_path)
    pyDecoder = io_utils.get_decoder(pyCfg)
    pySequence = importlib.import_module("reconstruct").get_sequence(seq_dir, pyCfg)
    # ---- src/LocalMapping.cc:38-40 ----
    optim = importlib.import_module("reconstruct.optimizer")
    pyOptimizer = optim.Optimizer(pyDecoder, pyCfg)
    pyMeshExtractor = optim.MeshExtractor(pyDecoder, pyCfg.optimizer.code_len, pyCfg.voxels_dim)
    assert pyOptimizer.code_len == 64                                   # LocalMapping_util.cc:416
    # ---- src/Tracking_util.cc:35-49 ----
    detections = pySequence.get_frame_by_id(0)
    assert pySequence.engine is pyOptimizer.engine                      # one packed decoder per process
    assert len(detections) >= 3
    n_good = 0
    for det in detections:
        pts = F(np.asarray(det.surface_points, np.float32))             # .cast<Eigen::MatrixXf>() and back
        Sim3Tco = F(np.asarray(det.T_cam_obj, np.float32))
        assert pts.ndim == 2 and pts.shape[1] == 3 and Sim3Tco.shape == (4, 4)
        if det.rays is None:
            continue
        rays, depth = F(np.asarray(det.rays, np.float32)), np.asarray(det.depth, np.float32)
        if pts.shape[0] < 50:                                           # LocalMapping_util.cc:170
            continue
        # ---- src/LocalMapping_util.cc:181-199 ----
        pyMapObject = pyOptimizer.reconstruct_object(Sim3Tco, pts, rays, depth)
        if not pyMapObject.is_good:
            assert pyMapObject.t_cam_obj is None and pyMapObject.code is None
            continue
        n_good += 1
        Two, code = pyMapObject.t_cam_obj, pyMapObject.code
        assert Two.shape == (4, 4) and Two.dtype == np.float32 and code.shape == (64,) and code.dtype == np.float32
        assert np.isfinite(float(pyMapObject.loss))
        pyMesh = pyMeshExtractor.extract_mesh_from_code(code)
        assert pyMesh.vertices.dtype == np.float32 and pyMesh.vertices.shape[1] == 3
        assert pyMesh.faces.dtype == np.int32 and pyMesh.faces.shape[1] == 3 and pyMesh.faces.shape[0] > 0
        # ---- src/LocalMapping_util.cc:94-111 (re-observation) ----
        s = float(np.cbrt(np.linalg.det(Two[:3, :3].astype(np.float64))))
        SE3Tco = Two.copy()
        SE3Tco[:3, :3] /= s
        T2 = pyOptimizer.estimate_pose_cam_obj(F(SE3Tco), s, pts, code)
        assert np.asarray(T2, np.float32).shape == (4, 4) and np.isfinite(np.asarray(T2)).all()
        # ---- src/MapObject_util.cc:43-46 ----
        pts_obj = (np.linalg.inv(Two.astype(np.float64)) @ np.c_[pts, np.ones(len(pts))].T).T[:, :3]
        v = float(pyOptimizer.compute_sdf_loss_objectpoint_zhjd(F(pts_obj.astype(np.float32)), code))
        assert abs(v) < 0.1
    assert n_good >= 2

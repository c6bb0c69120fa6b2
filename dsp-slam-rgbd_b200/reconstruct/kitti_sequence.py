"""Drop-in for the hot part of the reference's ``reconstruct/kitti_sequence.py``: the per-frame association of
3-D detections, LiDAR points and 2-D masks that produces the arguments of ``Optimizer.reconstruct_object``.

Same class and attribute names as the reference (`FrameWithLiDAR` :27-216, `KITIISequence` :219-277), so the C++
caller's ``pySequence.get_frame_by_id(frame_id)`` (`src/Tracking_util.cc:35-49`) and `reconstruct_frame.py:42`
keep working.  The numpy of `get_detections` runs in libdspslam_sm100.so (`dspslam/frame.py`, `csrc/frame.cu`),
bit-identical to the reference.  Out of scope and therefore NOT here: the online detectors (mmdetection) and the
image decoder - a frame is built from arrays (`velo_pts`, `detections_3d`, `det_2d`) or from the reference's
pre-stored label files (`configs.path_label_3d/2d`, torch-pickled) and velodyne ``.bin`` scans.
"""
from __future__ import annotations

import os

import numpy as np

from dspslam.frame import FrameBuilder
from reconstruct.utils import ForceKeyErrorDict  # noqa: F401  (instances are ForceKeyErrorDicts)


def read_calib_file(filepath):
    """KITTI `calib.txt` -> {key: float64 array}; parsing stops at the first blank line and entries that are not
    numeric (dates) are dropped, as in `reconstruct/utils.py:60-74`."""
    out = {}
    with open(filepath) as fh:
        for raw in fh:
            if raw == "\n":
                break
            name, _, rest = raw.partition(":")
            try:
                out[name] = np.array(rest.split(), dtype=np.float64)
            except ValueError:
                continue
    return out


def load_velo_scan(file):
    """(n, 4) float32 velodyne scan (`reconstruct/utils.py:77-80`)."""
    return np.fromfile(file, dtype=np.float32).reshape((-1, 4))


class FrameWithLiDAR:
    def __init__(self, sequence, frame_id, velo_pts=None, detections_3d=None, det_2d=None, img_hw=None):
        self.configs = sequence.configs
        self.K = sequence.K_cam
        self.invK = sequence.invK_cam
        self.T_cam_velo = sequence.T_cam_velo
        self.max_lidar_pts = self.configs.num_lidar_max
        self.min_lidar_pts = self.configs.num_lidar_min
        self.min_mask_area = self.configs.min_mask_area
        self.frame_id = frame_id
        self.sequence = sequence
        if velo_pts is None:
            velo_pts = load_velo_scan(os.path.join(sequence.velo_dir, "%06d.bin" % frame_id))
        if detections_3d is None or det_2d is None:
            import torch
            if detections_3d is None:
                detections_3d = torch.load(os.path.join(sequence.lbl3d_dir, "%06d.lbl" % frame_id), weights_only=False)
            if det_2d is None:
                det_2d = torch.load(os.path.join(sequence.lbl2d_dir, "%06d.lbl" % frame_id), weights_only=False)
        self.velo_pts = velo_pts
        self.detections_3d = np.asarray(detections_3d, np.float32).reshape(-1, 7)
        self.masks_2d = np.asarray(det_2d["pred_masks"])
        self.bboxes_2d = np.asarray(det_2d["pred_boxes"], np.float32)
        if img_hw is None:
            img_hw = self.masks_2d.shape[1:3] if self.masks_2d.ndim == 3 and self.masks_2d.shape[0] else (375, 1242)
        self.img_h, self.img_w = int(img_hw[0]), int(img_hw[1])
        self.instances = []
        self.batch = None        # dspslam.frame.FrameBatch: the same detections resident in HBM

    def get_detections(self):
        if self.detections_3d.shape[0] == 0:
            return
        builder = FrameBuilder(self.sequence.engine, self.K, self.invK, self.T_cam_velo, self.img_w, self.img_h,
                               self.max_lidar_pts, self.min_mask_area, self.configs.downsample_ratio)
        self.batch = builder.build_kitti(self.velo_pts, self.detections_3d, self.masks_2d, self.bboxes_2d)
        self.instances = builder.instances(self.batch, self.masks_2d, self.bboxes_2d)


class KITIISequence:
    def __init__(self, data_dir, configs, engine=None, decoder=None):
        """Reference signature `KITIISequence(data_dir, configs)` (`kitti_sequence.py:219-238`, called through
        `reconstruct.get_sequence` from `src/System.cc:99`).  Optional: `engine` (a dspslam.engine.Engine) or the
        `decoder` module the Optimizer uses; with neither, the engine is resolved at the first frame - the one the
        process's Optimizer already created, else one built from `get_decoder(configs)` (cached per DeepSDF_DIR, so
        it is the module `src/System.cc:98` holds)."""
        self.root_dir = data_dir
        self.rgb_dir = os.path.join(data_dir, "image_2")
        self.velo_dir = os.path.join(data_dir, "velodyne")
        self.calib_file = os.path.join(data_dir, "calib.txt")
        self.load_calib()
        self.configs = configs
        self.num_frames = len(os.listdir(self.rgb_dir)) if os.path.isdir(self.rgb_dir) else 0
        self.online = self.configs.detect_online
        self.lbl2d_dir = self.configs.path_label_2d
        self.lbl3d_dir = self.configs.path_label_3d
        if not self.online:
            assert self.lbl2d_dir is not None and self.lbl3d_dir is not None, "pre-stored label paths missing"
        from reconstruct import get_detectors
        self.detector_2d, self.detector_3d = get_detectors(self.configs)
        self._engine, self._decoder = engine, decoder
        self.current_frame = None
        self.detections_in_current_frame = None

    @property
    def engine(self):
        if self._engine is None:
            from reconstruct.optimizer import _resolve_engine
            self._engine = _resolve_engine(self._decoder, self.configs)
        return self._engine

    def load_calib(self):
        """Intrinsics of camera 2 and the velodyne -> camera-2 transform, FP32 (`kitti_sequence.py:243-257`):
        K = P2[:, :3]; camera 2 sits P2[0,3]/P2[0,0] along x from camera 0, whose pose in the velodyne frame is Tr."""
        cal = read_calib_file(self.calib_file)
        P2 = cal["P2"].reshape(3, 4)
        self.K_cam = P2[:, :3].astype(np.float32)
        self.invK_cam = np.linalg.inv(self.K_cam).astype(np.float32)
        cam0_from_velo = np.vstack([cal["Tr"].reshape(3, 4), [0.0, 0.0, 0.0, 1.0]])
        cam2_from_cam0 = np.eye(4)
        cam2_from_cam0[0, 3] = P2[0, 3] / P2[0, 0]
        self.T_cam_velo = cam2_from_cam0.dot(cam0_from_velo).astype(np.float32)

    def get_frame_by_id(self, frame_id, **arrays):
        self.current_frame = FrameWithLiDAR(self, frame_id, **arrays)
        self.current_frame.get_detections()
        self.detections_in_current_frame = self.current_frame.instances
        return self.detections_in_current_frame

"""Drop-in for the hot part of the reference's ``reconstruct/mono_sequence.py`` (`Frame.get_detections`
:80-114): background rays of the largest instance mask (lattice sampler, sub-sampling, lens undistortion, rays)
computed by libdspslam_sm100.so, bit-identical to the reference's numpy + cv2.undistortPoints.  Detectors and image
decoding are out of scope: a frame is built from the 2-D detection arrays or the pre-stored label file."""
from __future__ import annotations

import os

import numpy as np

from dspslam.frame import FrameBuilder
from reconstruct.utils import ForceKeyErrorDict


class Frame:
    def __init__(self, sequence, frame_id, det_2d=None, img_hw=None):
        self.configs = sequence.configs
        self.K = sequence.K_cam
        self.invK = sequence.invK_cam
        self.k1 = sequence.k1
        self.k2 = sequence.k2
        self.min_mask_area = self.configs.min_mask_area
        self.frame_id = frame_id
        self.sequence = sequence
        if det_2d is None:
            import torch
            det_2d = torch.load(os.path.join(sequence.lbl2d_dir, "%06d.lbl" % frame_id), weights_only=False)
        self.masks_2d = np.asarray(det_2d["pred_masks"])
        self.bboxes_2d = np.asarray(det_2d["pred_boxes"], np.float32)
        if img_hw is None and self.masks_2d.ndim == 3:
            img_hw = self.masks_2d.shape[1:3]
        self.img_h, self.img_w = int(img_hw[0]), int(img_hw[1])
        self.instances = []

    def get_detections(self):
        if self.masks_2d.shape[0] == 0:
            return
        builder = FrameBuilder(self.sequence.engine, self.K, self.invK, None, self.img_w, self.img_h,
                               min_mask_area=self.min_mask_area, downsample_ratio=self.configs.downsample_ratio,
                               k1=self.k1, k2=self.k2)
        max_id, rays = builder.background_rays_mono(self.masks_2d, self.bboxes_2d)
        instance = ForceKeyErrorDict()
        instance.bbox = self.bboxes_2d[max_id, ...]
        instance.mask = self.masks_2d[max_id, ...].astype(np.float32) * 255.
        instance.background_rays = rays
        self.instances = [instance]


def read_camera_yaml(path):
    """Camera.fx/fy/cx/cy/k1/k2 of an ORB-SLAM settings file (`mono_sequence.py:123-129` reads them with
    `cv2.FileStorage(...).getNode(name).real()`).  cv2 when importable, else the flat ``key: value`` lines of the
    OpenCV YAML dialect (``%YAML:1.0`` header, ``#`` comments) - the only structure those six keys use.  A missing
    key reads 0.0, as `FileNode.real()` of an empty node does."""
    names = ("Camera.fx", "Camera.fy", "Camera.cx", "Camera.cy", "Camera.k1", "Camera.k2")
    try:
        import cv2
        fs = cv2.FileStorage(path, cv2.FILE_STORAGE_READ)
        if not fs.isOpened():
            raise IOError("cannot open " + path)
        vals = [fs.getNode(n).real() for n in names]
        fs.release()
        return dict(zip(names, vals))
    except ImportError:
        vals = dict.fromkeys(names, 0.0)
        with open(path) as fh:
            for raw in fh:
                line = raw.split("#", 1)[0].strip()
                key, sep, val = line.partition(":")
                if sep and key.strip() in vals:
                    try:
                        vals[key.strip()] = float(val)
                    except ValueError:
                        pass
        return vals


class MonoSequence:
    def __init__(self, data_dir, configs, K_cam=None, k1=None, k2=None, engine=None, decoder=None):
        """Reference signature `MonoSequence(data_dir, configs)` (`mono_sequence.py:117-147`, called through
        `reconstruct.get_sequence` from `src/System.cc:99`): intrinsics and distortion come from the ORB-SLAM
        settings file `configs.slam_config_path`.  `K_cam, k1, k2` may be given instead (tests, tools); the engine is
        resolved as in `KITIISequence`."""
        self.root_dir = data_dir
        self.rgb_dir = os.path.join(data_dir, "image_0")
        if K_cam is None:
            cam = read_camera_yaml(configs.slam_config_path)
            K_cam = np.array([[cam["Camera.fx"], 0.0, cam["Camera.cx"]], [0.0, cam["Camera.fy"], cam["Camera.cy"]],
                              [0.0, 0.0, 1.0]])
            k1 = cam["Camera.k1"] if k1 is None else k1
            k2 = cam["Camera.k2"] if k2 is None else k2
        # float64 K as the reference keeps it (:130-131); the frame builder casts where the reference's numpy does
        self.K_cam = np.asarray(K_cam, np.float64).reshape(3, 3)
        self.invK_cam = np.linalg.inv(self.K_cam)
        self.k1, self.k2 = float(k1 or 0.0), float(k2 or 0.0)
        self.configs = configs
        self.data_type = self.configs.data_type
        assert self.data_type == "Redwood" or self.data_type == "Freiburg", "Wrong data type, supported: Redwood and Freiburg"
        self.online = self.configs.detect_online
        self.lbl2d_dir = self.configs.path_label_2d
        if not self.online:
            assert self.lbl2d_dir is not None, "pre-stored label path missing"
        from reconstruct import get_detectors
        self.detector_2d = get_detectors(self.configs)
        self._engine, self._decoder = engine, decoder
        self.current_frame = None
        self.detections_in_current_frame = None

    @property
    def engine(self):
        if self._engine is None:
            from reconstruct.optimizer import _resolve_engine
            self._engine = _resolve_engine(self._decoder, self.configs)
        return self._engine

    def get_frame_by_id(self, frame_id, **arrays):
        self.current_frame = Frame(self, frame_id, **arrays)
        self.current_frame.get_detections()
        self.detections_in_current_frame = self.current_frame.instances
        return self.detections_in_current_frame

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


class MonoSequence:
    def __init__(self, data_dir, configs, K_cam, k1=0.0, k2=0.0, engine=None, decoder=None):
        """The reference reads K, k1, k2 from the ORB-SLAM yaml through cv2.FileStorage (:124-136); here they are
        passed in (no OpenCV dependency)."""
        self.root_dir = data_dir
        self.rgb_dir = os.path.join(data_dir, "image_0")
        self.K_cam = np.asarray(K_cam, np.float32).reshape(3, 3)
        self.invK_cam = np.linalg.inv(self.K_cam).astype(np.float32)
        self.k1, self.k2 = float(k1), float(k2)
        self.configs = configs
        self.data_type = self.configs.data_type
        self.online = False
        self.lbl2d_dir = self.configs.path_label_2d
        if engine is None:
            from reconstruct.optimizer import _engine_for
            engine = _engine_for(decoder)
        self.engine = engine
        self.current_frame = None
        self.detections_in_current_frame = None

    def get_frame_by_id(self, frame_id, **arrays):
        self.current_frame = Frame(self, frame_id, **arrays)
        self.current_frame.get_detections()
        self.detections_in_current_frame = self.current_frame.instances
        return self.detections_in_current_frame

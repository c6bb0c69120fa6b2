"""Golden vectors for the detection -> optimiser-input builder.
TEST INFRASTRUCTURE: runs in the build container only (needs /root/reference); output: tests/golden/frame_*.npz.

The UNMODIFIED reference classes `reconstruct.kitti_sequence.FrameWithLiDAR` (kitti_sequence.py:27-216) and
`reconstruct.mono_sequence.MonoFrame`-style sampler (mono_sequence.py:51-114) are executed on synthetic frames
written to a scratch directory in the formats they read (velodyne .bin, .png, torch-saved .lbl label files).
Harness-side shims only: stub `addict`/`plyfile`/`skimage`, an empty `get_detectors`, `np.bool`/`np.bool8`
aliases (removed in numpy 2; the reference pins numpy 1.19).

    python oracle/make_frame_goldens.py
"""
from __future__ import annotations

import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
OUT = os.path.join(ROOT, "tests", "golden")

sys.path.insert(0, os.path.join(ROOT, "dsp-slam-rgbd_b200", "dspslam"))
from frame_synth import IMG_H, IMG_W, K, T_CAM_VELO, synth_kitti_frame  # noqa: E402  (numpy only; no package import:
# the reference's own `reconstruct` package must be the one imported in this process)


# MonoSequence builds FP64 intrinsics from the yaml (mono_sequence.py:123-131)
K_MONO = np.array([[535.4, 0.0, 620.5], [0.0, 539.2, 187.3], [0.0, 0.0, 1.0]])


def _shim_reference():
    import ref_harness
    ref_harness.import_reference()
    if not hasattr(np, "bool"):
        np.bool = bool            # kitti_sequence.py:178 (numpy 1.19 alias)
    if not hasattr(np, "bool8"):
        np.bool8 = np.bool_       # mono_sequence.py:99
    import reconstruct
    if not hasattr(reconstruct, "get_detectors"):
        reconstruct.get_detectors = lambda configs: (None, None)
    import reconstruct.kitti_sequence as ks
    import reconstruct.mono_sequence as ms
    return ks, ms


def run_reference_kitti(ks, scan, boxes, masks, bbs, cfg, scratch, frame_id=0):
    import cv2
    import torch
    for d in ("image_2", "velodyne", "l2d", "l3d"):
        os.makedirs(os.path.join(scratch, d), exist_ok=True)
    H, W = masks.shape[1:]
    cv2.imwrite(os.path.join(scratch, "image_2", "%06d.png" % frame_id), np.zeros((H, W, 3), np.uint8))
    scan.tofile(os.path.join(scratch, "velodyne", "%06d.bin" % frame_id))
    torch.save(boxes, os.path.join(scratch, "l3d", "%06d.lbl" % frame_id))
    torch.save({"pred_masks": masks, "pred_boxes": bbs}, os.path.join(scratch, "l2d", "%06d.lbl" % frame_id))
    _orig_load = torch.load
    torch.load = lambda f, *a, **k: _orig_load(f, *a, weights_only=False, **k)   # label files are pickled numpy
    seq = types.SimpleNamespace(
        configs=cfg, rgb_dir=os.path.join(scratch, "image_2"), velo_dir=os.path.join(scratch, "velodyne"),
        lbl2d_dir=os.path.join(scratch, "l2d"), lbl3d_dir=os.path.join(scratch, "l3d"), K_cam=K,
        invK_cam=np.linalg.inv(K).astype(np.float32), T_cam_velo=T_CAM_VELO, online=False,
        detector_2d=None, detector_3d=None)
    try:
        frame = ks.FrameWithLiDAR(seq, frame_id)
        frame.get_detections()
    finally:
        torch.load = _orig_load
    return frame.instances


def pack_instances(instances, masks):
    out = {"n_inst": np.int32(len(instances))}
    for i, ins in enumerate(instances):
        out["i%d_T_cam_obj" % i] = np.asarray(ins["T_cam_obj"], np.float32)
        out["i%d_scale" % i] = np.asarray(ins["scale"], np.float32)
        out["i%d_pts" % i] = np.asarray(ins["surface_points"], np.float32)
        out["i%d_is_front" % i] = np.bool_(ins["is_front"])
        has = ins["rays"] is not None
        out["i%d_has_rays" % i] = np.bool_(has)
        if has:
            out["i%d_rays" % i] = np.asarray(ins["rays"], np.float32)
            out["i%d_depth" % i] = np.asarray(ins["depth"], np.float32)
        if "mask" in ins:
            n = [k for k in range(masks.shape[0]) if np.array_equal(masks[k], ins["mask"])]
            out["i%d_mask_id" % i] = np.int32(n[0])
        else:
            out["i%d_mask_id" % i] = np.int32(-1)
    return out


def run_reference_mono(ms, masks, bbs, cfg, k1, k2, scratch, frame_id=0):
    import cv2
    import torch
    for d in ("image_0", "l2d"):
        os.makedirs(os.path.join(scratch, d), exist_ok=True)
    H, W = masks.shape[1:]
    cv2.imwrite(os.path.join(scratch, "image_0", "%06d.png" % frame_id), np.zeros((H, W, 3), np.uint8))
    torch.save({"pred_masks": masks, "pred_boxes": bbs}, os.path.join(scratch, "l2d", "%06d.lbl" % frame_id))
    _orig_load = torch.load
    torch.load = lambda f, *a, **k: _orig_load(f, *a, weights_only=False, **k)
    seq = types.SimpleNamespace(configs=cfg, rgb_dir=os.path.join(scratch, "image_0"),
                                lbl2d_dir=os.path.join(scratch, "l2d"), K_cam=K_MONO,
                                invK_cam=np.linalg.inv(K_MONO), k1=k1, k2=k2, online=False,
                                detector_2d=None, data_type="Redwood")
    try:
        frame = ms.MonoFrame(seq, frame_id) if hasattr(ms, "MonoFrame") else ms.Frame(seq, frame_id)
        frame.get_detections()
    finally:
        torch.load = _orig_load
    return frame.instances


def main():
    ks, ms = _shim_reference()
    from ref_harness import _Dict
    os.makedirs(OUT, exist_ok=True)
    cfg = _Dict(num_lidar_max=250, num_lidar_min=10, min_mask_area=1000, downsample_ratio=4.0)
    for seed, kw in ((0, {}), (1, dict(n_cars=8, n_ground=90000, pts_per_car=2500)), (2, dict(n_cars=6, n_ground=12000, pts_per_car=400))):
        scan, boxes, masks, bbs = synth_kitti_frame(seed, **kw)
        with tempfile.TemporaryDirectory() as scratch:
            inst = run_reference_kitti(ks, scan, boxes, masks, bbs, cfg, scratch)
        out = pack_instances(inst, masks)
        out.update(seed=np.int32(seed), kw=np.array(repr(kw)), K=K, T_cam_velo=T_CAM_VELO,
                   num_lidar_max=np.int32(250), min_mask_area=np.int32(1000), downsample_ratio=np.float32(4.0))
        path = os.path.join(OUT, "frame_kitti_%d.npz" % seed)
        np.savez_compressed(path, **out)
        print(path, "instances", len(inst), "with rays", sum(i["rays"] is not None for i in inst),
              "pts", [i["surface_points"].shape[0] for i in inst],
              "rays", [None if i["rays"] is None else i["rays"].shape[0] for i in inst])
    # monocular sampler (largest mask; undistorted background rays)
    mcfg = _Dict(min_mask_area=1000, downsample_ratio=4.0)
    for seed in (0, 1):
        _, _, masks, bbs = synth_kitti_frame(10 + seed, n_cars=3, n_ground=10)
        with tempfile.TemporaryDirectory() as scratch:
            inst = run_reference_mono(ms, masks, bbs, mcfg, -0.12, 0.05, scratch)
        path = os.path.join(OUT, "frame_mono_%d.npz" % seed)
        np.savez_compressed(path, seed=np.int32(10 + seed), K=K_MONO, k1=np.float64(-0.12), k2=np.float64(0.05),
                            bbox=np.asarray(inst[0]["bbox"], np.float32),
                            mask_id=np.int32(int(np.argmax(masks.sum(-1).sum(-1)))),
                            background_rays=np.asarray(inst[0]["background_rays"], np.float32))
        print(path, "bg rays", inst[0]["background_rays"].shape)


if __name__ == "__main__":
    main()

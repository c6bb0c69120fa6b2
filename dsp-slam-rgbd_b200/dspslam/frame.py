"""Detection -> optimiser-input builder on the GPU.

`FrameBuilder.build_kitti` turns one stereo+LiDAR frame (scan, 3-D boxes, 2-D instance masks) into the packed
batch `dspslam_gn_joint` reads - surface points, rays, depths, offsets - WITHOUT leaving HBM, so that
`Optimizer.reconstruct_frame` runs detection association and the Gauss-Newton loop back to back on the device.
It replaces the numpy of `reconstruct/kitti_sequence.py:112-216` (and `mono_sequence.py:80-114` for the
monocular background rays) bit for bit; only the per-detection 4x4 bookkeeping (`box_frames`, ~20 flops per
detection) stays on the host, written as the reference writes it.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N
from .engine import JointBatch, _align

F = np.float32


def box_frames(detections_3d, T_cam_velo):
    """Depth-ordered boxes -> (order, dspslam_frame_box array, T_cam_obj (B,4,4), size (B,3)).
    kitti_sequence.py:125-135 (depth order, T_velo_obj, inverse), :139-147 (3 m cube, half extents with w, l
    enlarged by 1.1), :156-157,163 (T_cam_obj scaled by l, is_front)."""
    det = np.asarray(detections_3d, F).reshape(-1, 7)
    order = np.argsort(det[:, 0])
    B = det.shape[0]
    boxes = (N.FrameBox * max(B, 1))()
    T_co_all = np.zeros((B, 4, 4), F)
    sizes = np.zeros((B, 3), F)
    T_cam_velo = np.asarray(T_cam_velo, F)
    for i, d in enumerate(det[order]):
        trans, size, theta = d[:3], d[3:6], d[6]
        T_vo = np.array([[np.cos(theta), 0, -np.sin(theta), trans[0]],
                         [-np.sin(theta), 0, -np.cos(theta), trans[1]],
                         [0, 1, 0, trans[2] + size[2] / 2],
                         [0, 0, 0, 1]]).astype(F)
        T_ov = np.linalg.inv(T_vo)
        r = 3.0
        w, l, h = list(size / 2)
        w *= 1.1
        l *= 1.1
        T_co = T_cam_velo @ T_vo
        T_co[:3, :3] *= l
        b = boxes[i]
        b.T_obj_velo[:] = [float(v) for v in np.asarray(T_ov[:3], F).reshape(-1)]
        b.lo[:] = [float(F(trans[k] - r)) for k in range(3)]
        b.hi[:] = [float(F(trans[k] + r)) for k in range(3)]
        b.half[:] = [float(F(w)), float(F(h)), float(F(l))]
        b.is_front = int(T_co[2, 3] > 0.0)
        T_co_all[i] = T_co
        sizes[i] = size
    return order, boxes, T_co_all, sizes


class FrameBatch(JointBatch):
    """A frame's detections resident in HBM in the layout of `Engine.run_joint`, plus the host-side summary."""
    info = None          # (B, FRAME_INFO) int32: n_in_box, M, n_in_fov, mask_id, has_rays, n_bg, hits, 0
    T_cam_obj = None     # (B, 4, 4) float32 initial Sim(3), depth order
    sizes = None         # (B, 3) float32 box sizes (instance.scale)
    order = None         # depth order (indices into the caller's detections_3d)
    pts_off = ray_off = None


class FrameBuilder:
    def __init__(self, engine, K, invK=None, T_cam_velo=None, img_w=1242, img_h=375, num_lidar_max=250,
                 min_mask_area=1000, downsample_ratio=4.0, max_background=200, k1=0.0, k2=0.0):
        self.engine = engine
        K64 = np.asarray(K, np.float64).reshape(3, 3)       # mono keeps FP64 intrinsics, KITTI FP32 ones
        invK64 = np.linalg.inv(K64) if invK is None else np.asarray(invK, np.float64).reshape(3, 3)
        K = np.asarray(K, F).reshape(3, 3)
        invK = np.linalg.inv(K).astype(F) if invK is None else np.asarray(invK, F).reshape(3, 3)
        self.T_cam_velo = np.eye(4, dtype=F) if T_cam_velo is None else np.asarray(T_cam_velo, F).reshape(4, 4)
        p = N.FrameParams()
        p.img_w, p.img_h = int(img_w), int(img_h)
        p.num_lidar_max, p.min_mask_area = int(num_lidar_max), int(min_mask_area)
        p.downsample_ratio, p.max_background = int(downsample_ratio), int(max_background)
        p.K[:] = [float(v) for v in K.reshape(-1)]
        p.invK[:] = [float(v) for v in invK.reshape(-1)]
        p.T_cam_velo[:] = [float(v) for v in self.T_cam_velo[:3].reshape(-1)]
        p.k1, p.k2 = float(k1), float(k2)
        p.K64[:] = [float(v) for v in K64.reshape(-1)]
        p.invK64[:] = [float(v) for v in invK64.reshape(-1)]
        self.prm = p

    # ---- stereo + LiDAR ----------------------------------------------------------------------------------------
    def build_kitti(self, velo_pts, detections_3d, masks_2d, bboxes_2d, codes=None):
        """One frame -> FrameBatch (device resident).  `velo_pts` (n,4) f32, `detections_3d` (B,7) f32,
        `masks_2d` (n2,H,W) bool/uint8, `bboxes_2d` (n2,4) f32; numpy arrays or device tensors."""
        eng, t = self.engine, self.engine.torch
        dev = eng.device
        order, boxes, T_co, sizes = box_frames(detections_3d, self.T_cam_velo)
        B = T_co.shape[0]
        if B == 0:
            raise ValueError("no detections")
        L = eng.code_len
        cap_p, cap_r = B * self.prm.num_lidar_max, B * (self.prm.num_lidar_max + self.prm.max_background)
        offs, cur = {}, 0
        for name, nbytes in (("pts_off", 4 * (B + 1)), ("ray_off", 4 * (B + 1)), ("n_fg", 4 * B),
                             ("info", 4 * N.FRAME_INFO * B), ("pts", 12 * cap_p), ("rays", 12 * cap_r),
                             ("depth", 4 * cap_r), ("poses0", 64 * B), ("codes0", 4 * L * B),
                             ("poses", 64 * B), ("codes", 4 * L * B), ("status", 4 * B), ("loss", 4 * B)):
            offs[name] = cur
            cur = _align(cur + nbytes)
        offs["io_end"] = cur
        host = t.empty(cur, dtype=t.uint8, pin_memory=True)
        blob = t.empty(cur, dtype=t.uint8, device=dev)
        hb = host.numpy()
        hb[offs["poses0"]: offs["poses0"] + 64 * B] = T_co.reshape(-1).view(np.uint8)
        c0 = np.zeros((B, L), F) if codes is None else np.ascontiguousarray(np.asarray(codes, F)[order][:, :L])
        hb[offs["codes0"]: offs["codes0"] + 4 * L * B] = c0.reshape(-1).view(np.uint8)
        lo, hi = offs["poses0"], offs["poses"]
        blob[lo:hi].copy_(host[lo:hi], non_blocking=True)

        velo = self._dev(velo_pts, t.float32).reshape(-1, 4)
        masks = self._dev(masks_2d, t.uint8)
        n_masks = int(masks.shape[0]) if masks.dim() == 3 else 0
        if n_masks and (masks.shape[1] != self.prm.img_h or masks.shape[2] != self.prm.img_w):
            raise ValueError("mask size does not match the frame size")
        bbs = self._dev(bboxes_2d, t.float32).reshape(-1, 4) if n_masks else None
        base = blob.data_ptr()
        N.check(eng.lib.dspslam_frame_build_kitti(
            eng._h, eng._stream(), C.byref(self.prm), velo.data_ptr(), velo.shape[0], boxes, B,
            masks.data_ptr() if n_masks else None, n_masks, bbs.data_ptr() if n_masks else None,
            base + offs["pts_off"], base + offs["ray_off"], base + offs["n_fg"], base + offs["pts"],
            base + offs["rays"], base + offs["depth"], base + offs["info"]), "dspslam_frame_build_kitti")
        # the only host round trip of the frame: (3B + 2 + 8B) int32 of offsets and counts
        lo, hi = offs["pts_off"], offs["pts"]
        host[lo:hi].copy_(blob[lo:hi], non_blocking=True)
        t.cuda.current_stream(dev).synchronize()
        pts_off = hb[offs["pts_off"]: offs["pts_off"] + 4 * (B + 1)].view(np.int32).copy()
        ray_off = hb[offs["ray_off"]: offs["ray_off"] + 4 * (B + 1)].view(np.int32).copy()
        info = hb[offs["info"]: offs["info"] + 4 * N.FRAME_INFO * B].view(np.int32).reshape(B, N.FRAME_INFO).copy()
        fb = FrameBatch(B, L, blob, host, offs, int(pts_off[-1]), int(ray_off[-1]),
                        int(np.diff(pts_off).max()), int(np.diff(ray_off).max()))
        fb.info, fb.T_cam_obj, fb.sizes, fb.order, fb.pts_off, fb.ray_off = info, T_co, sizes, order, pts_off, ray_off
        fb._keep = (velo, masks, bbs)
        return fb

    def instances(self, fb, masks_2d=None, bboxes_2d=None):
        """Host copies in the shape of the reference's `FrameWithLiDAR.instances` (kitti_sequence.py:159-209):
        T_cam_obj, scale, surface_points, num_surface_points, is_front, rays, depth [, mask, bbox]."""
        from reconstruct.utils import ForceKeyErrorDict
        t = self.engine.torch
        o = fb.offs
        lo, hi = o["pts"], o["poses0"]
        fb.host[lo:hi].copy_(fb.blob[lo:hi], non_blocking=True)
        t.cuda.current_stream(self.engine.device).synchronize()
        hb = fb.host.numpy()
        pts = hb[o["pts"]: o["pts"] + 12 * fb.total_pts].view(F).reshape(-1, 3)
        rays = hb[o["rays"]: o["rays"] + 12 * fb.total_rays].view(F).reshape(-1, 3)
        depth = hb[o["depth"]: o["depth"] + 4 * fb.total_rays].view(F)
        out = []
        for b in range(fb.B):
            ins = ForceKeyErrorDict()
            ins.T_cam_obj = fb.T_cam_obj[b].copy()
            ins.scale = fb.sizes[b].copy()
            ins.surface_points = pts[fb.pts_off[b]: fb.pts_off[b + 1]].copy()
            ins.num_surface_points = int(fb.pts_off[b + 1] - fb.pts_off[b])
            ins.is_front = bool(fb.T_cam_obj[b, 2, 3] > 0.0)
            ins.rays = None
            mid = int(fb.info[b, 3])
            if mid >= 0:
                ins.mask_id = mid
                if masks_2d is not None:
                    ins.mask = masks_2d[mid]
                if bboxes_2d is not None:
                    ins.bbox = bboxes_2d[mid]
            if fb.info[b, 4]:
                M = ins.num_surface_points
                ins.rays = rays[fb.ray_off[b]: fb.ray_off[b + 1]].copy()
                ins.depth = depth[fb.ray_off[b]: fb.ray_off[b] + M].copy()
            out.append(ins)
        return out

    # ---- monocular ---------------------------------------------------------------------------------------------
    def background_rays_mono(self, masks_2d, bboxes_2d):
        """mono_sequence.py:95-106: (mask_id, background_rays (n_bg,3) f32) of the largest mask."""
        eng, t = self.engine, self.engine.torch
        masks = self._dev(masks_2d, t.uint8)
        bbs = self._dev(bboxes_2d, t.float32).reshape(-1, 4)
        rays = t.empty((self.prm.max_background, 3), dtype=t.float32, device=eng.device)
        info = t.zeros(2, dtype=t.int32, device=eng.device)
        N.check(eng.lib.dspslam_frame_build_mono(eng._h, eng._stream(), C.byref(self.prm), masks.data_ptr(),
                                                 int(masks.shape[0]), bbs.data_ptr(), rays.data_ptr(),
                                                 info.data_ptr()), "dspslam_frame_build_mono")
        mid, n_bg = (int(v) for v in info.tolist())
        return mid, rays[:n_bg].cpu().numpy()

    def _dev(self, a, dtype):
        t = self.engine.torch
        if isinstance(a, t.Tensor):
            return a.to(device=self.engine.device, dtype=dtype).contiguous()
        a = np.asarray(a)
        if a.dtype == np.bool_:
            a = a.view(np.uint8)
        return t.from_numpy(np.ascontiguousarray(a)).to(device=self.engine.device, dtype=dtype)

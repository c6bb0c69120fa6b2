"""CPU oracle for the detection -> optimiser-input builder.
TEST INFRASTRUCTURE: imported only by tests/, never by the product path.

A numpy restatement of what the reference's sequence loaders compute between the detectors and
`Optimizer.reconstruct_object`:
  * `FrameWithLiDAR.get_detections`  reconstruct/kitti_sequence.py:112-216
  * `FrameWithLiDAR.pixels_sampler`  reconstruct/kitti_sequence.py:70-97  (same code in mono_sequence.py:51-78)
  * `get_rays`                       reconstruct/loss_utils.py:23-37
  * `Frame.get_detections` (mono)    reconstruct/mono_sequence.py:80-114, whose `cv2.undistortPoints` is a
    third-party step (opencv-python, `environment_cuda113.yml`): restated from OpenCV's published algorithm
    (`cvUndistortPointsInternal`: 5 fixed-point iterations of the radial/tangential model in double precision).

Parity is pinned: `tests/test_frame_builder.py` checks every function here against tests/golden/frame_*.npz,
which `oracle/make_frame_goldens.py` produced by executing the reference's own classes.

Arithmetic notes that the CUDA kernels reproduce bit for bit:
  * all point transforms are FP32 products summed left to right, `(a0 + a1) + a2`, then `+ t` (numpy's add-reduce
    over a length-3 axis), no fused multiply-add;
  * `np.linspace(a, b, n).astype(int32)` is `trunc(i * ((b - a) / (n - 1)) + a)` in FP64 with the last element set
    to `b` exactly;
  * rays are computed in FP64 (`concatenate` of FP32 pixel coordinates with int32 lattice pixels promotes to
    FP64, kitti_sequence.py:203-205) and rounded to FP32 once.
"""
from __future__ import annotations

import numpy as np

F = np.float32


def linspace_int(a, b, n):
    """`np.linspace(a, b, n).astype(np.int32)` spelled out (numpy/core/function_base.py)."""
    if n <= 0:
        return np.zeros(0, np.int32)
    if n == 1:
        return np.array([a], np.float64).astype(np.int32)
    step = (np.float64(b) - np.float64(a)) / np.float64(n - 1)
    y = np.arange(n, dtype=np.float64) * step + np.float64(a)
    y[-1] = b
    return y.astype(np.int32)


def xform_f32(R, t, p):
    """(p[:, None, :] * R).sum(-1) + t in FP32, left-to-right."""
    R, t, p = R.astype(F), t.astype(F), p.astype(F)
    o = np.empty((p.shape[0], 3), F)
    for i in range(3):
        o[:, i] = ((p[:, 0] * R[i, 0] + p[:, 1] * R[i, 1]) + p[:, 2] * R[i, 2]) + t[i]
    return o


def box_frames(det3d, T_cam_velo):
    """Per-detection 4x4 bookkeeping (kitti_sequence.py:125-135,156-157): depth order, T_velo_obj, its inverse,
    crop bounds, T_cam_obj.  Tiny host-side work in the product as well."""
    det3d = np.asarray(det3d, F)
    order = np.argsort(det3d[:, 0])
    out = []
    for d in det3d[order]:
        trans, size, th = d[:3], d[3:6], d[6]
        T_vo = np.array([[np.cos(th), 0, -np.sin(th), trans[0]], [-np.sin(th), 0, -np.cos(th), trans[1]],
                         [0, 1, 0, trans[2] + size[2] / 2], [0, 0, 0, 1]]).astype(F)
        T_ov = np.linalg.inv(T_vo)
        r = 3.0
        lo = np.array([trans[0] - r, trans[1] - r, trans[2] - r], F)
        hi = np.array([trans[0] + r, trans[1] + r, trans[2] + r], F)
        w, l, h = list(size / 2)
        w *= 1.1
        l *= 1.1
        T_co = T_cam_velo @ T_vo
        T_co[:3, :3] *= l
        out.append(dict(T_obj_velo=T_ov.astype(F), lo=lo, hi=hi, half=np.array([w, h, l], F), T_cam_obj=T_co.astype(F),
                        scale=size.astype(F)))
    return order, out


def lidar_surface_points(velo, box, T_cam_velo, num_lidar_max):
    """kitti_sequence.py:136-155: radius crop, oriented-box crop, linspace sub-sampling, camera frame."""
    p = velo[:, :3]
    near = np.all((p > box["lo"]) & (p < box["hi"]), axis=1)
    pn = p[near]
    po = xform_f32(box["T_obj_velo"][:3, :3], box["T_obj_velo"][:3, 3], pn)
    hf = box["half"]
    on = np.all((po > -hf) & (po < hf), axis=1)
    ps = pn[on]
    n = ps.shape[0]
    if n > num_lidar_max:
        ps = ps[linspace_int(0, n - 1, num_lidar_max)]
    return xform_f32(T_cam_velo[:3, :3], T_cam_velo[:3, 3], ps), n


def project(pts_cam, K):
    """kitti_sequence.py:185-189: FP32 pinhole projection, strict field-of-view test, truncation to pixels."""
    K = K.astype(F)
    h = np.empty_like(pts_cam)
    for i in range(3):
        h[:, i] = (pts_cam[:, 0] * K[i, 0] + pts_cam[:, 1] * K[i, 1]) + pts_cam[:, 2] * K[i, 2]
    uv = h[:, :2] / h[:, 2:3]
    return uv


def pixels_sampler(bbox, mask, img_w, img_h, alpha):
    """kitti_sequence.py:70-97: lattice over the expanded box (stride alpha), pixels NOT on the mask, row-major."""
    alpha = int(alpha)
    max_w, max_h = img_w - 1, img_h - 1
    l, t, r, b = [int(v) for v in np.asarray(bbox).astype(np.int32)]
    l = l - 5 if l > 5 else 0
    t = t - 5 if t > 5 else 0
    r = r + 5 if r < max_w - 5 else max_w
    b = b + 5 if b < max_h - 5 else max_h
    hh = linspace_int(t, b, int((b - t + 1) / alpha))
    ww = linspace_int(l, r, int((r - l + 1) / alpha))
    vv = np.repeat(hh, ww.shape[0])
    uu = np.tile(ww, hh.shape[0])
    keep = ~mask[vv, uu]
    return np.stack([uu[keep], vv[keep]], 1)


def get_rays_f64(pix, invK):
    """loss_utils.py:23-37 on FP64 pixel coordinates; FP32 result."""
    invK = np.asarray(invK).astype(np.float64)      # FP32 (KITTI) or FP64 (mono) intrinsics, promoted
    pix = pix.astype(np.float64)
    d = np.empty((pix.shape[0], 3), np.float64)
    for i in range(3):
        d[:, i] = (pix[:, 0] * invK[i, 0] + pix[:, 1] * invK[i, 1]) + 1.0 * invK[i, 2]
    return d.astype(F)


def kitti_frame(velo, det3d, masks, bboxes, K, invK, T_cam_velo, img_w, img_h, num_lidar_max=250,
                min_mask_area=1000, downsample_ratio=4.0, max_bg=200):
    """The whole of `get_detections`: list of dicts (T_cam_obj, scale, surface_points, is_front, mask_id, rays,
    depth) in depth order."""
    _, boxes = box_frames(det3d, T_cam_velo)
    inst = []
    for bx in boxes:
        pts, _ = lidar_surface_points(velo, bx, T_cam_velo, num_lidar_max)
        inst.append(dict(T_cam_obj=bx["T_cam_obj"], scale=bx["scale"], surface_points=pts,
                         is_front=bool(bx["T_cam_obj"][2, 3] > 0.0), mask_id=-1, rays=None, depth=None))
    if masks.shape[0] == 0:
        return inst
    for ins in inst:
        if not ins["is_front"]:
            continue
        pts = ins["surface_points"]
        uv = project(pts, K)
        fov = (uv[:, 0] > 0) & (uv[:, 0] < img_w) & (uv[:, 1] > 0) & (uv[:, 1] < img_h)
        px = uv[fov].astype(np.int32)
        hits = np.array([int(masks[n, px[:, 1], px[:, 0]].sum()) for n in range(masks.shape[0])])
        if hits.max() > px.shape[0] * 0.5:
            n = int(np.argmax(hits))
            ins["mask_id"] = n
            if int(masks[n].sum()) > min_mask_area:
                bg = pixels_sampler(bboxes[n], masks[n], img_w, img_h, downsample_ratio)
                if bg.shape[0] > max_bg:
                    bg = bg[linspace_int(0, bg.shape[0] - 1, max_bg)]
                ins["rays"] = np.concatenate([get_rays_f64(uv, invK), get_rays_f64(bg, invK)], 0)
                ins["depth"] = pts[:, 2].astype(F)
    return inst


def undistort_points(pix, K, k1, k2, iters=5):
    """cv2.undistortPoints(pix, K, (k1, k2, 0, 0, 0), P=K) (mono_sequence.py:103-104): FP32 in/out, FP64 inside."""
    fx, fy, cx, cy = (np.float64(K[0, 0]), np.float64(K[1, 1]), np.float64(K[0, 2]), np.float64(K[1, 2]))
    k1, k2 = np.float64(k1), np.float64(k2)
    pix = pix.astype(F).astype(np.float64)
    x0 = (pix[:, 0] - cx) / fx
    y0 = (pix[:, 1] - cy) / fy
    x, y = x0.copy(), y0.copy()
    for _ in range(iters):
        r2 = x * x + y * y
        icdist = 1.0 / (1.0 + (k2 * r2 + k1) * r2)
        bad = icdist < 0
        xn, yn = x0 * icdist, y0 * icdist
        x = np.where(bad, x0, xn)
        y = np.where(bad, y0, yn)
    return np.stack([x * fx + cx, y * fy + cy], 1).astype(F)


def mono_background_rays(masks, bboxes, K, invK, k1, k2, img_w, img_h, downsample_ratio=4.0, max_bg=200):
    """mono_sequence.py:95-106: largest mask, lattice sampler, sub-sampling, undistortion, rays."""
    n = int(np.argmax(masks.sum(axis=-1).sum(axis=-1)))
    bg = pixels_sampler(bboxes[n], masks[n], img_w, img_h, downsample_ratio)
    if bg.shape[0] > max_bg:
        bg = bg[linspace_int(0, bg.shape[0] - 1, max_bg)]
    und = undistort_points(bg, K, k1, k2)
    return n, get_rays_f64(und, invK)

"""Synthetic KITTI-style frames for the detection -> optimiser-input builder (numpy only, no package imports):
a LiDAR scan, 3-D boxes, instance masks and 2-D boxes in the formats the reference's sequence loaders read
(`reconstruct/kitti_sequence.py:99-216`).  Used by tests, tools and `oracle/make_frame_goldens.py`."""
import numpy as np

IMG_W, IMG_H = 1242, 375
K = np.array([[718.856, 0.0, 607.1928], [0.0, 718.856, 185.2157], [0.0, 0.0, 1.0]], np.float32)
T_CAM_VELO = np.array([[4.2768e-04, -9.9996e-01, -8.0845e-03, -1.1984e-02],
                       [-7.2106e-03, 8.0811e-03, -9.9994e-01, -5.4039e-02],
                       [9.9997e-01, 4.8594e-04, -7.2069e-03, -2.9219e-01],
                       [0, 0, 0, 1]], np.float32)


def synth_kitti_frame(seed, n_cars=5, n_ground=20000, pts_per_car=1500, img_wh=(IMG_W, IMG_H)):
    """LiDAR scan (N,4) f32 in the velodyne frame (x forward, y left, z up), 3-D boxes (n,7) f32
    [x, y, z_bottom, w, l, h, theta], instance masks (n2,H,W) bool and boxes (n2,4) f32 [l,t,r,b]."""
    rng = np.random.default_rng(seed)
    W, H = img_wh
    g = np.stack([rng.uniform(2, 60, n_ground), rng.uniform(-25, 25, n_ground),
                  -1.73 + 0.02 * rng.normal(size=n_ground)], 1)
    pts, boxes = [g], []
    for c in range(n_cars):
        x, y = rng.uniform(6, 40), rng.uniform(-7, 7)
        w, l, h = rng.uniform(1.5, 1.9), rng.uniform(3.4, 4.6), rng.uniform(1.4, 1.7)
        th = rng.uniform(-np.pi, np.pi)
        n = int(pts_per_car * rng.uniform(0.02, 1.0))          # some cars get fewer points than num_lidar_max
        # points on the faces of the box, object frame of the reference: x across (w), y up (h), z along (l)
        u = rng.uniform(-1, 1, (n, 3)) * np.array([w / 2, h / 2, l / 2])
        face = rng.integers(0, 3, n)
        sign = rng.choice([-1.0, 1.0], n)
        half = np.array([w / 2, h / 2, l / 2])
        u[np.arange(n), face] = sign * half[face]
        u += 0.01 * rng.normal(size=u.shape)
        T_velo_obj = np.array([[np.cos(th), 0, -np.sin(th), x], [-np.sin(th), 0, -np.cos(th), y],
                               [0, 1, 0, -1.73 + h / 2], [0, 0, 0, 1]])
        pv = u @ T_velo_obj[:3, :3].T + T_velo_obj[:3, 3]
        pts.append(pv)
        boxes.append([x, y, -1.73, w, l, h, th])
    # one detection behind the vehicle (is_front False) and one far away with almost no points
    boxes.append([-12.0, 2.0, -1.73, 1.7, 4.0, 1.5, 0.3])
    pts.append(np.stack([rng.uniform(-13, -11, 300), rng.uniform(1, 3, 300), rng.uniform(-1.7, -0.3, 300)], 1))
    scan = np.concatenate(pts).astype(np.float32)
    scan = scan[rng.permutation(scan.shape[0])]
    scan = np.concatenate([scan, rng.uniform(0, 1, (scan.shape[0], 1)).astype(np.float32)], 1)
    boxes = np.array(boxes, np.float32)
    # masks: filled convex-ish silhouettes = dilated projections of each car's box corners
    masks, bbs = [], []
    for b in boxes[:n_cars]:
        x, y, z0, w, l, h, th = [float(v) for v in b]
        cs = np.array([[sx * w / 2, sy * h / 2, sz * l / 2] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])
        T_velo_obj = np.array([[np.cos(th), 0, -np.sin(th), x], [-np.sin(th), 0, -np.cos(th), y],
                               [0, 1, 0, z0 + h / 2], [0, 0, 0, 1]])
        cv = cs @ T_velo_obj[:3, :3].T + T_velo_obj[:3, 3]
        cc = cv @ T_CAM_VELO[:3, :3].astype(np.float64).T + T_CAM_VELO[:3, 3]
        uv = (cc @ K.astype(np.float64).T)
        uv = uv[:, :2] / uv[:, 2:3]
        l_, t_, r_, b_ = uv[:, 0].min(), uv[:, 1].min(), uv[:, 0].max(), uv[:, 1].max()
        vv, uu = np.mgrid[0:H, 0:W]
        cx, cy, ax, ay = (l_ + r_) / 2, (t_ + b_) / 2, (r_ - l_) / 2 + 1, (b_ - t_) / 2 + 1
        m = ((uu - cx) / ax) ** 2 + ((vv - cy) / ay) ** 2 < 1.0
        masks.append(m)
        bbs.append([max(l_, 0), max(t_, 0), min(r_, W - 1), min(b_, H - 1)])
    # a distractor mask that matches nothing, and one tiny mask (area < min_mask_area) over car 0
    vv, uu = np.mgrid[0:H, 0:W]
    masks.append((uu < 60) & (vv < 40))
    bbs.append([0, 0, 59, 39])
    return scan, boxes, np.array(masks), np.array(bbs, np.float32)

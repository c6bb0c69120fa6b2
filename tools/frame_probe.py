"""A few frame-builder calls on a KITTI-sized synthetic frame (for `ncu` launch lists of csrc/frame.cu)."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [os.path.join(ROOT, "dsp-slam-rgbd_b200"), ROOT]
import torch  # noqa: E402
from dspslam import synth  # noqa: E402
from dspslam.engine import Engine  # noqa: E402
from dspslam.frame import FrameBuilder  # noqa: E402
from dspslam.frame_synth import IMG_H, IMG_W, K, T_CAM_VELO, synth_kitti_frame  # noqa: E402

scan, boxes, masks, bbs = synth_kitti_frame(3, n_cars=11, n_ground=110000, pts_per_car=1200)
eng = Engine(synth.load_wfit())
fb = FrameBuilder(eng, K, np.linalg.inv(K).astype(np.float32), T_CAM_VELO, IMG_W, IMG_H)
d = (torch.from_numpy(scan).cuda(), boxes, torch.from_numpy(masks.view(np.uint8)).cuda(), torch.from_numpy(bbs).cuda())
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    out = fb.build_kitti(*d)
torch.cuda.synchronize()
print("frame:", out.total_pts, "surface points,", out.total_rays, "rays")

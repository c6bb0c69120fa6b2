"""The stock-PyTorch port that `tools/bench_configs.py torchgn` times on the B200 (tools/torch_gn.py) must be the same
algorithm as the oracle, otherwise its objects/s is not a like-for-like number: one and two Gauss-Newton steps on a
C1-shaped detection, CPU, FP32, against `oracle/dspslam_oracle.py` (itself pinned to the reference's vectors)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.mark.parametrize("iters", [1, 2])
def test_torch_port_matches_oracle(iters):
    import torch_gn
    from dspslam import synth
    from oracle.dspslam_oracle import OracleDecoder, OracleOptimizer
    net = synth.load_wfit().eval()
    d = synth.make_kitti_detection(0, M=120, n_bg=60)
    cfg = synth.kitti_config()
    cfg.optimizer.joint_optim.num_iterations = iters
    T, z, good, loss = torch_gn.reconstruct_object(net, cfg, d["t_cam_obj"], d["pts"], d["rays"], d["depth"])
    o = OracleOptimizer(OracleDecoder(net.state_dict()), cfg).reconstruct_object(d["t_cam_obj"], d["pts"], d["rays"],
                                                                                   d["depth"])
    assert good and o["is_good"]
    dt, rot, _ = synth.pose_errors(T.numpy(), o["t_cam_obj"])
    assert dt < 1e-4 and rot < 1e-2                      # metres / degrees: FP32 summation-order noise only
    assert np.linalg.norm(z.numpy() - o["code"]) <= 1e-3 * max(1e-6, np.linalg.norm(o["code"]))
    assert abs(loss - float(o["loss"])) < 1e-4 * max(1.0, abs(float(o["loss"])))

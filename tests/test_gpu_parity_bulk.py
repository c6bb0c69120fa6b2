"""north_star's population statement, asserted on the GPU box: ">= 95 % of reference objects within tolerance" of
the UNMODIFIED reference (vectors: tests/golden/bulk_c1.npz - 128 KITTI-style cars x 10 iterations, bulk_c3.npz -
32 monocular detections x 20 iterations, queries_wfit.npz; made by oracle/make_goldens*.py from /root/reference).

The statistic is tools/parity_stats.py (the same function bench.py prints as `parity`): one Gauss-Newton step from
the reference's own state at iteration e against the reference's state at e + 1, within 1 cm / 0.1 deg / 1e-2
relative code; and the per-entry Jacobian error against `get_batch_sdf_jacobian`.

  fp32      FFMA validation mode: >= 99 % of the steps, every Jacobian entry within 1e-5
  tc        the benchmarked tensor-core mode (FP16 operands + FP32-grade second evaluation next to the decision
            thresholds + hi/lo weights in the SDF+Jacobian kernel): >= 95 % of the steps, >= 95 % of the objects
            (last step), >= 95 % of the Jacobian rows with every entry within 1e-3
  tc_fast   single FP16 pass everywhere (round-1 arithmetic): reported, only held to a floor
"""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.fixture(scope="module")
def eng():
    from dspslam import synth
    from dspslam.engine import Engine
    return Engine(synth.load_wfit())


@pytest.fixture(scope="module")
def stats(eng):
    import parity_stats
    cache = {}

    def get(mode):
        if mode not in cache:
            cache[mode] = parity_stats.parity(eng, mode)
            print(mode, cache[mode])
        return cache[mode]
    return get


def test_fp32_mode_population(stats):
    p = stats("fp32")
    assert p["jacobian"]["rows_within"] == 1.0 and p["jacobian"]["sdf_max_err"] < 1e-5
    assert p["c1"]["objects"] >= 100
    assert p["c1"]["steps_within"] >= 0.99 and p["c1"]["objects_last_step_within"] >= 0.99
    assert p["c3"]["steps_within"] >= 0.99
    assert p["c2"]["steps_within"] >= 0.99


def test_tensor_core_mode_population(stats):
    p = stats("tc")
    assert p["jacobian"]["sdf_max_err"] < 1e-3
    assert p["jacobian"]["rows_within"] >= 0.95 and p["jacobian"]["entries_within"] >= 0.998
    assert p["c1"]["objects"] >= 100
    assert p["c1"]["steps_within"] >= 0.95, p["c1"]
    assert p["c1"]["objects_last_step_within"] >= 0.95, p["c1"]
    assert p["c3"]["steps_within"] >= 0.95 and p["c3"]["objects_last_step_within"] >= 0.95, p["c3"]
    # pose alone is never the limiting criterion: every step of every object
    assert p["c1"]["dt_m"]["max"] < 1e-2 and p["c1"]["rot_deg"]["max"] < 0.1
    # the benchmarked C2 batch itself (16 detections x 10 iterations, M = 2048 / N = 2248): measured 159 of 160 steps
    assert p["c2"]["objects"] == 16 and p["c2"]["steps_within"] >= 0.97 and p["c2"]["objects_last_step_within"] == 1.0


def test_exact_tensor_core_mode_population(stats):
    """DSPSLAM_MODE_TC_EXACT: the SDF+Jacobian rows with both operands as hi+lo FP16 pairs in both sweeps (k_tc_jacx).
    Jacobian rows of the 600 reference queries with every entry within 1e-3: 599-600 (one row has a ReLU unit within
    1e-6 of its kink, where ~22-bit operands can still flip it; typical entry error 3e-6, SDF 2.5e-6), >= 99 % of the
    GN steps and of the objects within the north_star tolerances."""
    p = stats("tc_exact")
    assert p["jacobian"]["rows_within"] >= 0.995 and p["jacobian"]["entries_within"] >= 0.9999
    assert p["jacobian"]["sdf_max_err"] < 2e-5
    assert p["c1"]["steps_within"] >= 0.99 and p["c1"]["objects_last_step_within"] >= 0.98, p["c1"]
    assert p["c3"]["steps_within"] >= 0.99, p["c3"]
    assert p["c2"]["steps_within"] >= 0.99 and p["c2"]["objects_last_step_within"] == 1.0, p["c2"]


def test_fast_mode_is_the_less_exact_one(stats):
    """The second evaluation and the hi/lo weights are what buys the parity: without them the same statistic drops
    (measured 0.88 of the steps); the fast mode is only held to a floor and must not beat the default."""
    fast, tc = stats("tc_fast"), stats("tc")
    assert fast["c1"]["steps_within"] >= 0.80
    assert fast["c1"]["steps_within"] <= tc["c1"]["steps_within"]
    assert fast["c1"]["code_rel"]["q95"] > tc["c1"]["code_rel"]["q95"]


def test_jacobian_outlier_rows_are_relu_kink_rows(eng):
    """The Jacobian rows of the tensor-core mode that are NOT within 1e-3 of the reference in every entry (2-3 % of the
    rows, 0.08 % of the entries) are counted and attributed: each of them has a hidden unit whose FP32 pre-activation
    lies within 1e-4 of zero (the FP16 rounding scale of the activations), i.e. a ReLU that the two evaluations
    take on different sides - there the derivative itself jumps.  Rows without such a unit are all within 1e-3."""
    import numpy as np
    from dspslam import synth
    from oracle import dspslam_oracle as orc
    with np.load(os.path.join(ROOT, "tests", "golden", "queries_wfit.npz")) as z:
        g = {k: z[k] for k in z.files}
    _, jac = eng.sdf_jac(g["code"], g["xyz"], mode="tc")
    err = np.abs(jac.cpu().numpy() - g["jac"])
    bad = err.max(axis=1) > 1e-3
    od = orc.OracleDecoder(synth.load_wfit().state_dict())
    inp = np.concatenate([np.broadcast_to(g["code"], (g["xyz"].shape[0], 64)), g["xyz"]], -1).astype(np.float32)
    h, mins = inp, []
    for k in range(od.n_layers - 1):
        if k in od.latent_in:
            h = np.concatenate([h, inp], -1)
        pre = h @ od.W[k].T + od.b[k]
        mins.append(np.abs(pre).min(axis=1))
        h = np.maximum(pre, 0)
    near_kink = np.min(np.stack(mins), axis=0) < 1e-4
    assert bad.mean() < 0.05
    assert np.all(near_kink[bad]), "a Jacobian row off by more than 1e-3 without a ReLU unit at its kink"
    assert err[~near_kink].max() <= 1e-3 if (~near_kink).any() else True
    assert err.max() < 0.2

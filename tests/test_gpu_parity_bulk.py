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

Two further populations (round 3):
  held-out  bulk_c1_heldout.npz: 128 cars with other seeds (oracle/make_goldens_heldout.py) that played no part in
            choosing the default hi/lo layer mask - the >= 95 % statement is asserted on them as well
  free run  north_star's end-state sentence as worded: all 10 iterations free-running, END state against the
            reference's.  The reference does not meet it against ITSELF (tests/golden/noise_floor_bulk.json: surface
            points moved by one FP32 ulp -> 51 % of 128 objects within tolerance), so the yardstick is that floor.
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


def _floor():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "noise_floor_bulk.json")) as f:
        return json.load(f)["summary"]


def test_held_out_population(stats):
    """The teacher-forced statistic on 128 detections that were NOT used to pick the default hi/lo layer mask
    (api.cu: wsplit_mask).  Measured: fp32 0.999 / 1.0, tc 0.9555 of 1 280 steps / 0.961 of the objects (last step),
    tc_exact 0.991 / 0.977.  Binomial standard error at n = 1 280 is 0.6 points: the default mode meets the 95 %
    statement on fresh data with no margin to spare, and the docs say so."""
    f, t, x = stats("fp32")["c1_heldout"], stats("tc")["c1_heldout"], stats("tc_exact")["c1_heldout"]
    assert f["objects"] == t["objects"] == 128
    assert f["steps_within"] >= 0.99 and f["objects_last_step_within"] >= 0.99
    assert t["steps_within"] >= 0.95 and t["objects_last_step_within"] >= 0.95, t
    assert x["steps_within"] >= 0.985 and x["objects_last_step_within"] >= 0.97, x


def test_free_run_end_state_against_the_reference_floor(stats):
    """Free-running end state (north_star's wording) of every shipped mode next to the reference's own
    reproducibility.  Fractions of 128 objects within 1 cm / 0.1 deg / 1e-2, bulk_c1 | held-out:
        reference vs itself (+1 ulp)   0.51
        fp32                           0.59 | 0.51      at the floor
        tc_exact (window 1.0)          0.47 | 0.38      within ~1-3 binomial sigma (4.4 points) of the floor
        tc (benchmarked default)       0.16 | 0.06      KNOWN GAP: FP16 Jacobian operands, see DESIGN.md section 5
    Asserted: the FP32 mode reaches the floor minus 5 points; the exact mode stays within 15 points of it and well
    above the default; the default mode is only guarded against getting worse (loose criterion 5 cm / 1 deg / 0.1)."""
    floor = _floor()
    assert floor["objects"] == 128 and 0.3 < floor["within"] < 0.8
    for key in ("c1_free_run", "c1_heldout_free_run"):
        f, x, t = stats("fp32")[key], stats("tc_exact")[key], stats("tc")[key]
        assert f["within"] >= floor["within"] - 0.05, (key, f)
        assert f["within_5cm_1deg_1e-1"] >= 0.90, (key, f)
        assert x["within"] >= floor["within"] - 0.15, (key, x)
        assert x["within"] >= t["within"] + 0.15, (key, x, t)
        assert x["within_5cm_1deg_1e-1"] >= 0.85, (key, x)
        assert t["within_5cm_1deg_1e-1"] >= 0.65 and t["dt_m"]["median"] < 3e-3 and t["rot_deg"]["median"] < 0.2, (key, t)

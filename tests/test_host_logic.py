"""Host-side logic that needs no GPU: decoder-variant checks, the slab schedule of the sharded grid, the verbatim copy of
the reference used by the reference arm, the bench's shared config."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_piece_ranges_cover_the_volume_exactly_once():
    from dspslam.dist import piece_ranges
    for total in (1, 7, 64, 125, 729, 16777216):
        for world in (1, 2, 3, 8):
            for pieces in (1, 3, 4):
                seen = []
                per = None
                for rank in range(world):
                    per, rs = piece_ranges(total, rank, world, pieces)
                    for j, (lo, hi, out_lo) in enumerate(rs):
                        assert 0 <= lo <= hi <= total and hi - lo <= per
                        assert out_lo == j * world * per
                        if hi > lo:
                            assert lo == out_lo + rank * per          # position in the gathered buffer == global index
                            seen.append((lo, hi))
                seen.sort()
                assert seen[0][0] == 0 and seen[-1][1] == total
                assert all(a[1] == b[0] for a, b in zip(seen[:-1], seen[1:]))


def test_unsupported_decoder_variants_are_rejected():
    from dspslam import synth
    from dspslam.packer import check_variant
    dec = synth.load_wfit()
    check_variant(dec, dec.state_dict())
    for attr, val in (("use_tanh", True), ("xyz_in_all", True), ("latent_in", (2, 4))):
        old = getattr(dec, attr, None)
        setattr(dec, attr, val)
        with pytest.raises(ValueError):
            check_variant(dec, dec.state_dict())
        setattr(dec, attr, old)
    with pytest.raises(ValueError):
        check_variant(None, {"lin0.bias": 0, "bn0.weight": 0, "bn0.bias": 0})


def test_oracle_ref_is_a_verbatim_copy_of_the_reference():
    """oracle/_ref (what the reference arm of bench.py times on the GPU box) holds the reference's own files,
    unmodified: sha256 of every copied file against /root/reference (skipped where the reference is absent)."""
    ref, dst = "/root/reference", os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref, "reconstruct")):
        pytest.skip("/root/reference is not on this machine")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fetch_ref
    assert fetch_ref.fetch()
    with open(os.path.join(dst, "MANIFEST.json")) as f:
        man = json.load(f)["sha256"]
    assert set(man) == set(fetch_ref.FILES)
    for rel, digest in man.items():
        for root in (ref, dst):
            with open(os.path.join(root, rel), "rb") as f:
                assert hashlib.sha256(f.read()).hexdigest() == digest, (root, rel)


def test_bench_arms_share_one_config():
    sys.path.insert(0, ROOT)
    import bench
    a, b = bench.shared_config(), bench.shared_config("tc", 0)
    assert a == b and "C2" in a["workload"] and a["objects_per_gpu"] == 16
    assert bench.traffic_of("k_tc<3>") is None or bench.traffic_of("k_tc<3>") > 1e6


def test_staging_ring_alternates_without_reallocating():
    """Engine._staging hands out two pinned buffers per size class alternately (no CUDA needed for the bookkeeping)."""
    import torch
    from dspslam.engine import Engine
    e = Engine.__new__(Engine)
    e.torch, e._pinned = torch, {}
    real_empty = torch.empty
    calls = []

    def fake_empty(n, dtype=None, pin_memory=False):
        calls.append(n)
        return real_empty(n, dtype=dtype)
    torch.empty = fake_empty
    try:
        a = e._staging(5000)
        b = e._staging(6000)
        c = e._staging(5000)
    finally:
        torch.empty = real_empty
    assert calls == [8192, 8192]                       # one pair for the 8 KiB class, allocated together
    assert a.numel() == 5000 and b.numel() == 6000
    assert a.data_ptr() != b.data_ptr() and a.data_ptr() == c.data_ptr()

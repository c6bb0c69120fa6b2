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


def test_staging_ring_is_bounded_in_bytes(monkeypatch):
    """The pinned pairs an engine keeps stay under PINNED_CAP_BYTES: least recently used classes go first, and a class
    that cannot fit on its own is handed out without being kept."""
    import torch
    from dspslam import engine as E
    e = E.Engine.__new__(E.Engine)
    e.torch, e._pinned = torch, {}
    real_empty = torch.empty
    monkeypatch.setattr(torch, "empty", lambda n, dtype=None, pin_memory=False: real_empty(n, dtype=dtype))
    monkeypatch.setattr(E, "PINNED_CAP_BYTES", 64 << 10)
    e._staging(5000)                 # 8 KiB class: 16 KiB kept
    e._staging(12000)                # 16 KiB class: 32 KiB kept (48 KiB total)
    e._staging(5000)                 # touches the 8 KiB class: the 16 KiB class is now the oldest
    e._staging(10000 + 8192)         # 32 KiB class = 64 KiB on its own: everything older has to go
    assert sorted(k[1] for k in e._pinned) == [32768]
    big = e._staging(40000)          # 64 KiB class: a pair would be 128 KiB > cap, not kept
    assert big.numel() == 40000 and sorted(k[1] for k in e._pinned) == [32768]


def test_last_engine_never_crosses_devices():
    """last_engine(device) only returns an engine living on that device (fake engines: no CUDA needed)."""
    import weakref
    import torch
    from dspslam import engine as E

    class Fake:
        def __init__(self, idx):
            self._h, self.device = 1, torch.device("cuda", idx)
    a, b = Fake(0), Fake(1)
    saved = list(E._LIVE)
    try:
        E._LIVE[:] = [weakref.ref(a), weakref.ref(b)]
        assert E.last_engine(0) is a and E.last_engine(1) is b and E.last_engine(2) is None
        b._h = None                  # closed engines are skipped
        assert E.last_engine(1) is None
    finally:
        E._LIVE[:] = saved


def test_wilson_interval_of_the_parity_fractions():
    """tools/parity_stats.py reports a 95 % interval next to each population fraction (the margins over 95 %
    are thin, so say how thin).  Known answers: 1 223 of 1 280 -> about [0.943, 0.966]."""
    import sys as _s
    _s.path.insert(0, os.path.join(ROOT, "tools"))
    import parity_stats
    lo, hi = parity_stats.wilson(1223, 1280)
    assert abs(lo - 0.9427) < 2e-3 and abs(hi - 0.9655) < 2e-3
    lo, hi = parity_stats.wilson(128, 128)
    assert hi == pytest.approx(1.0) and 0.96 < lo < 0.98
    assert parity_stats.wilson(0, 0) == [0.0, 1.0]

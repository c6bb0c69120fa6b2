"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/dspslam.h declares; the host-side mirror keeps the reference's names and error behaviour."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_library_exports_every_declared_symbol():
    from dspslam import _native
    if not os.path.isfile(_native.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    hdr = open(os.path.join(ROOT, "include", "dspslam.h")).read()
    declared = set(re.findall(r"\b(dspslam_[a-z0-9_]+)\s*\(", hdr))
    assert declared and declared == set(_native.EXPORTS)
    lib = ctypes.CDLL(_native.LIB_PATH)
    for sym in declared:
        assert hasattr(lib, sym), sym
    lib.dspslam_abi_version.restype = ctypes.c_int
    assert lib.dspslam_abi_version() == _native.ABI_VERSION


def test_product_path_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from dspslam import synth
    from dspslam._native import NativeError
    from reconstruct.optimizer import Optimizer
    with pytest.raises(NativeError):
        Optimizer(synth.make_decoder(0), synth.kitti_config())


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "dsp-slam-rgbd_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(d, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "/root/reference" not in src, f


def test_force_key_error_dict_and_configs(tmp_path):
    from reconstruct.utils import ForceKeyErrorDict, get_configs
    d = ForceKeyErrorDict(a=1, nested=dict(b=2))
    assert d.a == 1 and d.nested.b == 2 and d["nested"]["b"] == 2
    with pytest.raises(KeyError):
        d.missing
    with pytest.raises(KeyError):
        d.nested.missing
    p = tmp_path / "cfg.json"
    p.write_text('{"data_type": "KITTI", "optimizer": {"code_len": 64}}')
    cfg = get_configs(str(p))
    assert cfg.optimizer.code_len == 64
    with pytest.raises(KeyError):
        cfg.optimizer.joint_optim


def test_decoder_mirror_loads_deepsdf_checkpoint(tmp_path):
    import torch
    from deep_sdf.workspace import config_decoder
    from dspslam import synth
    from dspslam.packer import check_supported, effective_weights
    dec = synth.make_decoder(0)
    synth.write_experiment_dir(dec, str(tmp_path))
    dec2 = config_decoder(str(tmp_path))
    x = torch.randn(5, 67)
    assert torch.allclose(dec(x), dec2.cpu()(x))
    layers = effective_weights(dec.state_dict())
    assert check_supported(layers) == 64
    # folded weights reproduce the module's forward
    from oracle.dspslam_oracle import OracleDecoder
    od = OracleDecoder(dec.state_dict())
    assert np.abs(od.forward(x.numpy())[:, 0] - dec(x).detach().numpy()[:, 0]).max() < 1e-6
    for (w, _), wo in zip(layers, od.W):
        assert np.array_equal(w, wo)


def test_ply_writer(tmp_path):
    from reconstruct.utils import write_mesh_to_ply
    v = np.random.rand(4, 3).astype(np.float32)
    f = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    p = tmp_path / "m.ply"
    write_mesh_to_ply(v, f, str(p))
    raw = p.read_bytes()
    assert raw.startswith(b"ply\nformat binary_little_endian") and len(raw.split(b"end_header\n")[1]) == 4 * 12 + 2 * 13

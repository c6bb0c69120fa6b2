"""Host-side helpers of the reconstruction path (drop-in for the hot-path part
of the reference's ``reconstruct/utils.py``).

Mirrored names and what they follow in the reference:
  ForceKeyErrorDict        reconstruct/utils.py:82-84  (addict.Dict subclass whose
                           missing keys raise KeyError; re-implemented without addict)
  get_configs              reconstruct/utils.py:87-90
  get_decoder              reconstruct/utils.py:93-94
  create_voxel_grid        reconstruct/utils.py:97-116 (incl. the true-division
                           shear of modern torch)
  convert_sdf_voxels_to_mesh  reconstruct/utils.py:119-140 (marching cubes at level 0;
                           here the repo's own CUDA marching-cubes, see optimizer.py)
  write_mesh_to_ply        reconstruct/utils.py:143-163 (ASCII-free binary PLY writer,
                           no plyfile dependency)
"""
from __future__ import annotations

import json
import os

import numpy as np


class ForceKeyErrorDict(dict):
    """dict with attribute access; nested dicts are wrapped; a missing key
    raises ``KeyError`` both through ``d[k]`` and ``d.k``."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        for k, v in dict(*args, **kwargs).items():
            self[k] = v

    @classmethod
    def _wrap(cls, v):
        if isinstance(v, dict) and not isinstance(v, ForceKeyErrorDict):
            return cls(**v)
        if isinstance(v, (list, tuple)):
            return type(v)(cls._wrap(x) for x in v)
        return v

    def __setitem__(self, k, v):
        super().__setitem__(k, self._wrap(v))

    def __missing__(self, key):
        raise KeyError(key)

    def __getattr__(self, key):
        if key.startswith("__") and key.endswith("__"):
            raise AttributeError(key)
        return self[key]

    def __setattr__(self, key, value):
        self[key] = value

    def __delattr__(self, key):
        del self[key]


def get_configs(cfg_file):
    with open(cfg_file) as f:
        cfg_dict = json.load(f)
    return ForceKeyErrorDict(**cfg_dict)


_decoders = {}


def get_decoder(configs):
    """`reconstruct/utils.py:93-94`.  The module is cached per experiment directory: every caller in the process
    (System.cc:98, the sequence classes, scripts) then holds the SAME nn.Module, hence the same packed engine."""
    from deep_sdf.workspace import config_decoder
    key = os.path.abspath(configs.DeepSDF_DIR)
    if key not in _decoders:
        _decoders[key] = config_decoder(configs.DeepSDF_DIR)
    return _decoders[key]


def create_voxel_grid(vol_dim=128, coord_mode="reference_shear"):
    """(vol_dim^3, 3) float32 query grid over [-1, 1]^3, x slowest, z fastest.

    ``reference_shear`` reproduces what the reference computes under torch>=1.6
    (`utils.py:106-108`: ``index / vol_dim`` is a *true* division, so the y and x
    "indices" carry a fractional part); ``lattice`` is the regular grid the code
    was meant to build.  Host-side restatement used by tests and small tools; the
    CUDA grid kernel generates the same coordinates from the linear index.
    """
    n = vol_dim ** 3
    i = np.arange(n, dtype=np.int64)
    voxel_size = np.float32(2.0 / (vol_dim - 1))
    out = np.empty((n, 3), dtype=np.float32)
    zi = (i % vol_dim).astype(np.float32)
    if coord_mode == "reference_shear":
        fi = i.astype(np.float32)                       # int64 -> f32 (true_divide result dtype)
        q = fi / np.float32(vol_dim)
        yi = np.fmod(q, np.float32(vol_dim))
        xi = np.fmod(q / np.float32(vol_dim), np.float32(vol_dim))
    elif coord_mode == "lattice":
        yi = ((i // vol_dim) % vol_dim).astype(np.float32)
        xi = ((i // (vol_dim * vol_dim)) % vol_dim).astype(np.float32)
    else:
        raise ValueError("coord_mode must be 'reference_shear' or 'lattice'")
    out[:, 0] = xi * voxel_size + np.float32(-1)
    out[:, 1] = yi * voxel_size + np.float32(-1)
    out[:, 2] = zi * voxel_size + np.float32(-1)
    return out


def convert_sdf_voxels_to_mesh(pytorch_3d_sdf_tensor):
    """Same name, argument and return convention as the reference (`reconstruct/utils.py:119-140`): a (n, n, n) SDF
    volume (torch tensor, host or device, or numpy array) -> (vertices (nv, 3) float32 in [-1, 1]^3, faces (nf, 3)
    int32) at level 0 with voxel size 2 / (n - 1).  The reference runs skimage's `marching_cubes_lewiner` on the CPU;
    here the volume is meshed on the GPU by `dspslam_marching_cubes` (csrc/mesh.cu) through the process's engine -
    same vertex set, this repository's own triangulation table (dspslam/mesh.py; face topology against Lewiner's
    is unpinned: skimage is absent from the image)."""
    import torch
    from dspslam.mesh import sdf_grid_to_mesh
    from reconstruct.optimizer import _resolve_engine
    eng = _resolve_engine()
    sdf = pytorch_3d_sdf_tensor if isinstance(pytorch_3d_sdf_tensor, torch.Tensor) else torch.from_numpy(
        np.ascontiguousarray(pytorch_3d_sdf_tensor, dtype=np.float32))
    verts, faces = sdf_grid_to_mesh(eng, sdf.detach().to(eng.device))
    return verts, faces


def write_mesh_to_ply(v, f, ply_filename_out):
    """Binary little-endian PLY with float32 x,y,z vertices and int32 triangle lists."""
    v = np.ascontiguousarray(v, dtype="<f4")
    f = np.ascontiguousarray(f, dtype="<i4")
    header = ("ply\nformat binary_little_endian 1.0\n"
              "element vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
              "element face %d\nproperty list uchar int vertex_indices\nend_header\n"
              % (v.shape[0], f.shape[0]))
    rec = np.empty(f.shape[0], dtype=[("n", "u1"), ("idx", "<i4", (3,))])
    rec["n"] = 3
    rec["idx"] = f
    with open(ply_filename_out, "wb") as fh:
        fh.write(header.encode("ascii"))
        fh.write(v.tobytes())
        fh.write(rec.tobytes())

"""Marching cubes (csrc/mesh.cu): the generated case table and the numpy restatement on
CPU; the CUDA kernels against the restatement and against analytic / topological properties on the GPU."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def sphere_grid(G, r=0.6, c=(0.05, -0.02, 0.03)):
    ax = np.linspace(-1, 1, G, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    return (np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - r).astype(np.float32)


def smooth_random_grid(G, seed):
    rng = np.random.default_rng(seed)
    ax = np.linspace(-1, 1, G, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    f = np.zeros((G, G, G), np.float32)
    for _ in range(6):
        k = rng.uniform(1, 5, 3)
        f += rng.normal() * np.sin(k[0] * x + rng.uniform(0, 6)) * np.sin(k[1] * y + rng.uniform(0, 6)) * np.sin(k[2] * z + rng.uniform(0, 6))
    f += 100 * (np.maximum(np.abs(x), np.maximum(np.abs(y), np.abs(z))) - 0.8).clip(0)  # positive near the boundary: closed surfaces
    return f.astype(np.float32)


def test_table_header_is_generated():
    assert subprocess.call([sys.executable, os.path.join(ROOT, "tools", "gen_mc_table.py"), "--check"]) == 0


def test_table_cases_use_only_crossing_edges():
    from oracle import mc_oracle
    tab = mc_oracle.load_table()
    for case in range(256):
        ins = [(case >> k) & 1 for k in range(8)]
        used = set(int(e) for e in tab[case] if e >= 0)
        crossing = set()
        for e in range(12):
            a, j = divmod(e, 4)
            u, w = j & 1, j >> 1
            lo = [(0, u, w), (u, 0, w), (u, w, 0)][a]
            hi = list(lo)
            hi[a] = 1
            if ins[lo[0] + 2 * lo[1] + 4 * lo[2]] != ins[hi[0] + 2 * hi[1] + 4 * hi[2]]:
                crossing.add(e)
        assert used == crossing, case
        assert (tab[case] >= 0).sum() % 3 == 0


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_oracle_mesh_is_closed_oriented_manifold(seed):
    """random smooth fields hit the ambiguous cases; the face rule of the generator keeps the mesh watertight"""
    from oracle import mc_oracle
    v, f = mc_oracle.marching_cubes(smooth_random_grid(20, seed))
    assert len(f) > 100
    boundary, nonmanifold, conflicts, _ = mc_oracle.mesh_topology(f)
    assert boundary == 0 and nonmanifold == 0 and conflicts == 0


def test_oracle_sphere_geometry():
    from oracle import mc_oracle
    G, r = 24, 0.6
    v, f = mc_oracle.marching_cubes(sphere_grid(G, r))
    assert mc_oracle.mesh_topology(f) == (0, 0, 0, 2)
    c = np.array([0.05, -0.02, 0.03], np.float32)
    assert np.abs(np.linalg.norm(v - c, axis=1) - r).max() < 0.5 * (2.0 / (G - 1)) ** 2 / r + 1e-3
    a, b, cc = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    n = np.cross(b - a, cc - a)
    assert (np.einsum("ij,ij->i", n, (a + b + cc) / 3 - c) > 0).all()          # outward normals
    assert abs(0.5 * np.linalg.norm(n, axis=1).sum() / (4 * np.pi * r * r) - 1) < 0.02


def test_map_objects_txt_roundtrip(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "dsp-slam-rgbd_b200"))
    from dspslam import mesh
    T = np.eye(4)
    T[:3, :] = np.arange(12).reshape(3, 4) * 0.25
    code = np.linspace(-1, 1, 64).astype(np.float32)
    p = str(tmp_path / "MapObjects.txt")
    mesh.write_map_objects_txt(p, [(7, T, code), (9, np.eye(4), code * 2)])
    got = list(mesh.read_map_objects(p))
    assert [g[0] for g in got] == [7, 9]
    assert np.allclose(got[0][1], T) and np.allclose(got[0][2], code, atol=1e-7)


# ---------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def eng():
    from dspslam import synth
    from dspslam.engine import Engine
    return Engine(synth.load_wfit())


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["sphere", "rand0", "rand1"])
def test_gpu_marching_cubes_equals_oracle(which, eng):
    import torch
    from oracle import mc_oracle
    G = 19
    sdf = sphere_grid(G) if which == "sphere" else smooth_random_grid(G, int(which[-1]) + 10)
    v, f = eng.marching_cubes(torch.from_numpy(sdf))
    vo, fo = mc_oracle.marching_cubes(sdf)
    assert np.array_equal(f.cpu().numpy(), fo)                                 # indices: bit-exact
    assert np.allclose(v.cpu().numpy(), vo, atol=2e-6)                         # positions: FP32 rounding of one fma


@pytest.mark.gpu
def test_gpu_marching_cubes_capacity_retry_and_empty(eng):
    import torch
    sdf = torch.from_numpy(sphere_grid(96, r=0.9))                             # more vertices than the first guess? no: forces nothing
    v, f = eng.marching_cubes(sdf)
    assert v.shape[0] > 0 and f.shape[0] > 0 and int(f.max()) == v.shape[0] - 1
    from dspslam import _native as N
    counts = torch.zeros(3, dtype=torch.int32, device=v.device)
    N.check(eng.lib.dspslam_marching_cubes(eng._h, eng._stream(), sdf.to(v.device).data_ptr(), 96, 0.0, 2.0 / 95, -1.0,
                                           0, 0, None, None, counts.data_ptr()), "count only")
    assert counts.tolist() == [v.shape[0], f.shape[0], 0]                      # count-only call: does not fit, nothing written
    v0, f0 = eng.marching_cubes(torch.ones(8, 8, 8))
    assert v0.shape == (0, 3) and f0.shape == (0, 3)


@pytest.mark.gpu
def test_gpu_sphere_256_properties(eng):
    """BASELINE C4 size: closed oriented manifold, Euler characteristic 2, area within 0.1 %."""
    import torch
    from oracle import mc_oracle
    G, r = 256, 0.6
    v, f = eng.marching_cubes(torch.from_numpy(sphere_grid(G, r)))
    v, f = v.cpu().numpy(), f.cpu().numpy()
    assert mc_oracle.mesh_topology(f) == (0, 0, 0, 2)
    a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1).sum()
    assert abs(area / (4 * np.pi * r * r) - 1) < 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fp32", "tc"])
def test_mesh_extractor_end_to_end(mode, eng, tmp_path):
    """MeshExtractor.extract_mesh_from_code (optimizer.py:216-233): vertices lie on the decoder's zero level set."""
    from dspslam import synth, mesh
    from oracle import mc_oracle
    from reconstruct.optimizer import MeshExtractor
    from reconstruct.utils import write_mesh_to_ply
    code = np.zeros(64, np.float32)
    code[:3] = [0.3, -0.2, 0.1]
    ex = MeshExtractor(synth.load_wfit(), 64, 64, mode=mode, coord_mode="lattice")
    m = ex.extract_mesh_from_code(code)
    assert m.vertices.dtype == np.float32 and m.faces.dtype == np.int32 and m.vertices.shape[0] > 500
    boundary, nonmanifold, conflicts, _ = mc_oracle.mesh_topology(m.faces)
    assert boundary == 0 and nonmanifold == 0 and conflicts == 0
    s = eng.sdf(code, m.vertices, mode="fp32").cpu().numpy()
    assert np.abs(s).max() < 2e-2 and np.abs(s).mean() < 1e-3                  # linear interpolation across ReLU kinks, 64^3 grid
    write_mesh_to_ply(m.vertices, m.faces, str(tmp_path / "m.ply"))
    d = str(tmp_path)
    mesh.write_map_objects_txt(os.path.join(d, "MapObjects.txt"), [(1, np.eye(4), code)])
    assert mesh.extract_map_objects(d, ex) == [1]
    assert os.path.getsize(os.path.join(d, "objects", "1.ply")) > 1000


def crossing_set(sdf, level=0.0):
    """Closed form of the marching-cubes vertex SET: one vertex per grid edge whose end points lie on different
    sides of the level, linearly interpolated (the same set skimage's Lewiner variant produces; only the faces
    differ between case tables).  Vectorised numpy, any G."""
    sdf = np.asarray(sdf, np.float32)
    G = sdf.shape[0]
    sp = np.float32(2.0 / (G - 1))
    inside = sdf < np.float32(level)
    out = []
    for a in range(3):
        lo = [slice(None)] * 3
        hi = [slice(None)] * 3
        lo[a], hi[a] = slice(0, G - 1), slice(1, G)
        m = inside[tuple(lo)] != inside[tuple(hi)]
        idx = np.argwhere(m).astype(np.float32)
        v0, v1 = sdf[tuple(lo)][m], sdf[tuple(hi)][m]
        idx[:, a] += (np.float32(level) - v0) / (v1 - v0)
        out.append(idx * sp + np.float32(-1))
    return np.concatenate(out)


def mesh_volume_area(v, f):
    a, b, c = (v[f[:, k]].astype(np.float64) for k in range(3))
    vol = np.einsum("ij,ij->i", a, np.cross(b, c)).sum() / 6.0
    area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1).sum()
    return abs(vol), area


@pytest.mark.gpu
def test_c4_grid_mesh_vertex_set_volume_area(eng):
    """SURVEY 8(f) rank 1 at the BASELINE C4 size (decoder grid, G = 256): the GPU mesh's vertex set equals the
    closed-form crossing set; its enclosed volume agrees with the voxel count of the inside region; at G = 64 volume
    and area equal the numpy restatement's.  Face topology against skimage's `marching_cubes_lewiner`
    (reconstruct/utils.py:127-129) stays unpinned (skimage is not in the image)."""
    from oracle import mc_oracle
    code = np.zeros(64, np.float32)
    code[:3] = [0.2, -0.1, 0.3]
    for G in (64, 256):
        sdf = eng.grid_sdf(code, G, coord_mode="lattice", mode="tc").view(G, G, G)
        v, f = eng.marching_cubes(sdf)
        v, f, h = v.cpu().numpy(), f.cpu().numpy(), sdf.cpu().numpy()
        want = crossing_set(h)
        assert v.shape == want.shape
        from scipy.spatial import cKDTree
        # same count, and every vertex of either set has a partner in the other within FP32 rounding of the
        # interpolation (not asserted one-to-one: a level crossing exactly at a grid node is emitted by two edges)
        assert cKDTree(want).query(v)[0].max() < 4e-6 and cKDTree(v).query(want)[0].max() < 4e-6
        assert mc_oracle.mesh_topology(f)[:3] == (0, 0, 0)
        vol, area = mesh_volume_area(v, f)
        vox = float((h < 0).sum()) * (2.0 / (G - 1)) ** 3
        assert abs(vol / vox - 1) < (0.05 if G == 64 else 0.015)
        if G == 64:
            vo, fo = mc_oracle.marching_cubes(h)
            vol_o, area_o = mesh_volume_area(vo, fo)
            assert abs(vol / vol_o - 1) < 1e-5 and abs(area / area_o - 1) < 1e-5


@pytest.mark.gpu
def test_convert_sdf_voxels_to_mesh_and_cli(eng, tmp_path):
    """Same-named mirrors of the reference's mesh entry points: `convert_sdf_voxels_to_mesh` (utils.py:119-140) and
    the `extract_map_objects.py` command line (extract_map_objects.py:25-63)."""
    import json
    import torch
    from dspslam import mesh, synth
    from reconstruct.utils import convert_sdf_voxels_to_mesh
    sdf = torch.from_numpy(sphere_grid(40, r=0.5))
    v, f = convert_sdf_voxels_to_mesh(sdf)
    assert v.dtype == np.float32 and v.shape[1] == 3 and f.shape[1] == 3 and f.max() == v.shape[0] - 1
    assert np.abs(np.linalg.norm(v - np.array([0.05, -0.02, 0.03], np.float32), axis=1) - 0.5).max() < 2e-3
    d = str(tmp_path)
    synth.write_experiment_dir(synth.load_wfit(), os.path.join(d, "deepsdf"))
    cfg = dict(synth.kitti_config())
    cfg["DeepSDF_DIR"] = os.path.join(d, "deepsdf")
    with open(os.path.join(d, "config.json"), "w") as fh:
        json.dump(cfg, fh)
    code = np.zeros(64, np.float32)
    code[:3] = [0.3, -0.2, 0.1]
    mesh.write_map_objects_txt(os.path.join(d, "MapObjects.txt"), [(7, np.eye(4), code), (9, np.eye(4), -code)])
    env = dict(os.environ, PYTHONPATH=os.path.join(ROOT, "dsp-slam-rgbd_b200"))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "dsp-slam-rgbd_b200", "extract_map_objects.py"), "-c",
                        os.path.join(d, "config.json"), "-m", d, "-n", "48"], env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-1000:]
    for oid in (7, 9):
        assert os.path.getsize(os.path.join(d, "objects", "%d.ply" % oid)) > 1000
        assert np.load(os.path.join(d, "objects", "%d.npy" % oid)).shape == (4, 4)

"""Mesh extraction behind `MeshExtractor.extract_mesh_from_code` (reference: reconstruct/optimizer.py:216-233,
reconstruct/utils.py:119-163) and the map-object export of extract_map_objects.py:46-63.

The reference copies the (G,G,G) SDF volume to the host and runs skimage's Lewiner marching cubes on the CPU.
Here the volume stays on the GPU: `dspslam_marching_cubes` (csrc/mesh.cu) returns a welded, deterministically
ordered triangle mesh in device buffers; only vertices and faces cross to the host.  Conventions kept from the
reference: vertex = -1 + voxel_size * grid coordinate with voxel_size = 2 / (G - 1), level 0, vertices float32,
faces int32.  Differences (documented): the case table is this repository's own watertight
256-case table (tools/gen_mc_table.py), not Lewiner's 33-case one - the vertex SET (one per sign-changing grid
edge, linearly interpolated) is the same, the triangulation of ambiguous cells and the ordering differ; the
triangles are wound with normals towards larger SDF (outwards).
"""
from __future__ import annotations

import numpy as np

from reconstruct.utils import write_mesh_to_ply


def sdf_grid_to_mesh(engine, sdf, level=0.0):
    """(G,G,G) device tensor -> (vertices (nv,3) float32, faces (nf,3) int32) numpy arrays."""
    v, f = engine.marching_cubes(sdf, level=level)
    return v.cpu().numpy(), f.cpu().numpy()


def transform_vertices(vertices, t_world_obj):
    """object-frame vertices -> world frame with the Sim(3) pose of the map object (extract_map_objects.py:55-58)."""
    T = np.asarray(t_world_obj, np.float64)
    return (np.asarray(vertices, np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)


def read_map_objects(path_txt):
    """`MapObjects.txt` as written by the C++ side (src/System_util.cc:118-147): per object three lines -
    id, the 3x4 object pose (12 numbers), the shape code.  Yields (id, pose (4,4) float64, code float32)."""
    with open(path_txt) as fh:
        lines = fh.readlines()
    for i in range(len(lines) // 3):
        oid = int(lines[3 * i])
        pose = np.array([float(x) for x in lines[3 * i + 1].split()], np.float64).reshape(3, 4)
        pose = np.concatenate([pose, np.array([[0., 0., 0., 1.]])], axis=0)
        code = np.array([float(x) for x in lines[3 * i + 2].split()], np.float32)
        yield oid, pose, code


def write_map_objects_txt(path_txt, objects):
    """inverse of read_map_objects: `objects` = iterable of (id, pose (4,4), code)."""
    with open(path_txt, "w") as fh:
        for oid, T, code in objects:
            fh.write("%d\n" % int(oid))
            fh.write(" ".join("%.9g" % v for v in np.asarray(T, np.float64)[:3, :].reshape(-1)) + "\n")
            fh.write(" ".join("%.9g" % v for v in np.asarray(code, np.float64).reshape(-1)) + "\n")


def extract_map_objects(map_dir, extractor):
    """extract_map_objects.py:40-63: for every object of `map_dir`/MapObjects.txt write objects/<id>.npy (pose) and
    objects/<id>.ply (mesh of its shape code).  Returns the object ids."""
    import os
    save_dir = os.path.join(map_dir, "objects")
    os.makedirs(save_dir, exist_ok=True)
    ids = []
    for oid, pose, code in read_map_objects(os.path.join(map_dir, "MapObjects.txt")):
        np.save(os.path.join(save_dir, "%d.npy" % oid), pose)
        mesh = extractor.extract_mesh_from_code(code)
        write_mesh_to_ply(mesh.vertices, mesh.faces, os.path.join(save_dir, "%d.ply" % oid))
        ids.append(oid)
    return ids

"""TEST INFRASTRUCTURE: numpy restatement of csrc/mesh.cu (marching cubes with the generated 256-case table).

The reference's `skimage.measure.marching_cubes_lewiner` (reconstruct/utils.py:127-129, scikit-image 0.18.3) is not
installed in this image and its tables are not part of /root/reference, so mesh parity against the reference is
UNPINNED.  What is pinned here is the geometry both algorithms share - one vertex per sign-changing grid edge at the
linear interpolation of the level - plus topological properties (closed, consistently oriented 2-manifold)."""
import os
import re

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
TABLE_H = os.path.join(ROOT, "dsp-slam-rgbd_b200", "csrc", "mc_table.h")


def load_table():
    txt = open(TABLE_H).read()
    body = txt[txt.index("MC_TRI_HOST"):]
    rows = re.findall(r"\{([^{}]*)\}", body)
    tab = np.array([[int(v) for v in r.split(",")] for r in rows], np.int64)
    assert tab.shape == (256, 16)
    return tab


def marching_cubes(sdf, level=0.0, spacing=None, origin=-1.0):
    """-> vertices (nv,3) float32, faces (nf,3) int32 in the order mesh.cu produces."""
    sdf = np.asarray(sdf, np.float32)
    G = sdf.shape[0]
    spacing = np.float32(2.0 / (G - 1) if spacing is None else spacing)
    origin = np.float32(origin)
    level = np.float32(level)
    inside = sdf < level
    cross = np.zeros((G, G, G, 3), bool)
    cross[:-1, :, :, 0] = inside[:-1] != inside[1:]
    cross[:, :-1, :, 1] = inside[:, :-1] != inside[:, 1:]
    cross[:, :, :-1, 2] = inside[:, :, :-1] != inside[:, :, 1:]
    vid = (np.cumsum(cross.reshape(-1)) - 1).reshape(G, G, G, 3)
    nodes = np.argwhere(cross)                                    # (nv, 4): x, y, z, axis in emission order
    verts = np.zeros((nodes.shape[0], 3), np.float32)
    for n, (x, y, z, a) in enumerate(nodes):
        v0 = sdf[x, y, z]
        d = [0, 0, 0]
        d[a] = 1
        v1 = sdf[x + d[0], y + d[1], z + d[2]]
        t = np.float32(level - v0) / np.float32(v1 - v0)
        p = np.array([x, y, z], np.float32)
        p[a] = p[a] + t
        verts[n] = p * spacing + origin
    tab = load_table()
    faces = []
    c = np.zeros((G - 1, G - 1, G - 1), np.int64)
    for k in range(8):
        dx, dy, dz = k & 1, (k >> 1) & 1, (k >> 2) & 1
        c |= inside[dx:G - 1 + dx, dy:G - 1 + dy, dz:G - 1 + dz].astype(np.int64) << k
    for x, y, z in np.argwhere((c > 0) & (c < 255)):
        row = tab[c[x, y, z]]
        for t in range(0, 15, 3):
            if row[t] < 0:
                break
            tri = []
            for e in row[t:t + 3]:
                a, j = divmod(int(e), 4)
                u, w = j & 1, j >> 1
                off = [(0, u, w), (u, 0, w), (u, w, 0)][a]
                tri.append(vid[x + off[0], y + off[1], z + off[2], a])
            faces.append(tri)
    return verts, np.array(faces, np.int32).reshape(-1, 3)


def mesh_topology(faces):
    """(n_boundary_edges, n_nonmanifold_edges, n_orientation_conflicts, euler characteristic)"""
    f = np.asarray(faces, np.int64)
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    key = np.sort(e, axis=1)
    uniq, inv, cnt = np.unique(key, axis=0, return_inverse=True, return_counts=True)
    inv = inv.reshape(-1)
    sign = np.where(e[:, 0] < e[:, 1], 1, -1)
    bal = np.zeros(len(uniq), np.int64)
    np.add.at(bal, inv, sign)
    nv = len(np.unique(f))
    return int((cnt == 1).sum()), int((cnt > 2).sum()), int(((cnt == 2) & (bal != 0)).sum()), nv - len(uniq) + len(f)

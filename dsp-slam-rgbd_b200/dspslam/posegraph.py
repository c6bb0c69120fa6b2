"""Camera-object pose-graph edges on the device.

Host-side mirror of what the reference's bundle adjustment does with the measurement of `estimate_pose_cam_obj`:
`EdgeSE3LieAlgebra` (include/ObjectPoseGraph.h:56-88) between a keyframe vertex T_cw and an object vertex T_ow, as
built in src/Optimizer_util.cc:199-225 / :555-583 (information = invSigmaObject * I, Huber kernel) and read back
through `chi2()` (:647-656, :701-711).  The arithmetic runs in `csrc/posegraph.cu` through the C ABI
(`dspslam_pg_from_matrices`, `dspslam_pg_linearize`, `dspslam_pg_oplus`); there is no CPU fallback.

The sparse solve of the full bundle adjustment (map points, ORB features, g2o's Levenberg-Marquardt) stays where it
is in the reference; this module produces what that solver consumes from these edges, and `refine_objects` is the
closed sub-problem that needs nothing else: object vertices against FIXED keyframes, 6x6 systems per object.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N

# src/Optimizer_util.cc:82-84 (global BA) and :448-450 (local BA)
INV_SIGMA_OBJECT = 1e3
HUBER_GLOBAL_BA = float(np.sqrt(np.float32(0.10 * 1e3)))
HUBER_LOCAL_BA = float(np.sqrt(np.float32(1e3)))


class EdgeLinearization:
    """Device tensors of one `linearize` call (all FP64)."""
    __slots__ = ("error", "Ji", "Jj", "chi2", "rho", "weight", "A", "b", "Hij")

    def __init__(self, **kw):
        for k in self.__slots__:
            setattr(self, k, kw.get(k))


class PoseGraph:
    """Keyframe and object vertices plus camera-object edges, resident on the engine's device.

    vertices: (n, 7) g2o Vector7d rows (tx, ty, tz, qx, qy, qz, qw) - build them with `from_matrices` from the 4x4
    FP32 poses the SLAM system stores (`Converter::toSE3Quat`).  edge_i / edge_j index the keyframe / object vertex
    of every edge, measurements are the detections' `SE3Tco`.
    """

    def __init__(self, engine, inv_sigma=INV_SIGMA_OBJECT, huber_delta=HUBER_LOCAL_BA):
        self.engine = engine
        self.inv_sigma = float(inv_sigma)
        self.huber_delta = float(huber_delta)
        self.vertices = None
        self.edge_i = self.edge_j = self.measurements = None

    # -- construction -------------------------------------------------------------------------------------
    def from_matrices(self, T):
        """(n, 4, 4) FP32 poses -> (n, 7) FP64 device tensor."""
        eng, t = self.engine, self.engine.torch
        T = self._dev(T, t.float32).reshape(-1, 16)
        out = t.empty((T.shape[0], 7), dtype=t.float64, device=eng.device)
        N.check(eng.lib.dspslam_pg_from_matrices(eng._h, eng._stream(), T.data_ptr(), int(T.shape[0]),
                                                 out.data_ptr()), "dspslam_pg_from_matrices")
        return out

    def set_vertices(self, poses7):
        self.vertices = self._dev(poses7, self.engine.torch.float64).reshape(-1, 7).clone()

    def set_edges(self, edge_i, edge_j, measurements7):
        t = self.engine.torch
        self.edge_i = self._dev(edge_i, t.int32).reshape(-1)
        self.edge_j = self._dev(edge_j, t.int32).reshape(-1)
        self.measurements = self._dev(measurements7, t.float64).reshape(-1, 7)
        if not (self.edge_i.shape[0] == self.edge_j.shape[0] == self.measurements.shape[0]):
            raise ValueError("edge_i, edge_j and measurements must have one row per edge")
        if self.edge_i.numel() and self.vertices is not None:
            hi = int(max(self.edge_i.max(), self.edge_j.max()))
            lo = int(min(self.edge_i.min(), self.edge_j.min()))
            if lo < 0 or hi >= self.vertices.shape[0]:
                raise ValueError("edge refers to a vertex outside [0, %d)" % self.vertices.shape[0])

    # -- the edge class ----------------------------------------------------------------------------------
    def linearize(self, quadratic_form=True, robust=True):
        """computeError + linearizeOplus + chi2 (+ Huber) for every edge; with `quadratic_form` also the per-vertex
        blocks A, b and the per-edge off-diagonal blocks g2o's `constructQuadraticForm` adds up."""
        eng, t = self.engine, self.engine.torch
        nv, ne = int(self.vertices.shape[0]), int(self.edge_i.shape[0])
        f = dict(dtype=t.float64, device=eng.device)
        out = EdgeLinearization(error=t.empty((ne, 6), **f), Ji=t.empty((ne, 6, 6), **f), Jj=t.empty((ne, 6, 6), **f),
                                chi2=t.empty(ne, **f), rho=t.empty(ne, **f), weight=t.empty(ne, **f))
        if quadratic_form:
            out.A, out.b, out.Hij = t.empty((nv, 6, 6), **f), t.empty((nv, 6), **f), t.empty((ne, 6, 6), **f)
        p = lambda x: x.data_ptr() if x is not None else None    # noqa: E731
        N.check(eng.lib.dspslam_pg_linearize(
            eng._h, eng._stream(), self.vertices.data_ptr(), nv, self.edge_i.data_ptr(), self.edge_j.data_ptr(),
            self.measurements.data_ptr(), ne, C.c_double(self.inv_sigma),
            C.c_double(self.huber_delta if robust else 0.0), p(out.error), p(out.Ji), p(out.Jj), p(out.chi2),
            p(out.rho), p(out.weight), p(out.A), p(out.b), p(out.Hij)), "dspslam_pg_linearize")
        return out

    def chi2(self):
        """`e->chi2()` of every edge (the outlier test of Optimizer_util.cc:651,706 compares it with a threshold)."""
        return self.linearize(quadratic_form=False, robust=False).chi2

    def oplus(self, updates):
        """vertices <- exp(update) * vertices (VertexSE3Expmap::oplusImpl); updates (n, 6) = (omega, upsilon)."""
        eng, t = self.engine, self.engine.torch
        u = self._dev(updates, t.float64).reshape(-1, 6)
        if u.shape[0] != self.vertices.shape[0]:
            raise ValueError("one update row per vertex")
        N.check(eng.lib.dspslam_pg_oplus(eng._h, eng._stream(), self.vertices.data_ptr(), u.data_ptr(),
                                         int(u.shape[0])), "dspslam_pg_oplus")

    # -- closed sub-problem: objects against fixed keyframes ---------------------------------------------------
    def refine_objects(self, object_vertices, iterations=5, damping=0.0):
        """Gauss-Newton on the listed object vertices with every other vertex fixed: the object blocks of the
        reference's BA decouple then (each object sees only its own camera-object edges), so every step is one
        `linearize` and one batched 6x6 solve.  Returns the sum of rho over the edges before every step and after
        the last one."""
        t = self.engine.torch
        idx = self._dev(object_vertices, t.int64).reshape(-1)
        eye = t.eye(6, dtype=t.float64, device=self.engine.device)
        hist = []
        for _ in range(iterations):
            lin = self.linearize()
            hist.append(float(lin.rho.sum()))
            upd = t.zeros((self.vertices.shape[0], 6), dtype=t.float64, device=self.engine.device)
            upd[idx] = t.linalg.solve(lin.A[idx] + damping * eye, lin.b[idx].unsqueeze(-1)).squeeze(-1)
            self.oplus(upd)
        hist.append(float(self.linearize(quadratic_form=False).rho.sum()))
        return hist

    def _dev(self, a, dtype):
        t = self.engine.torch
        if isinstance(a, t.Tensor):
            return a.to(device=self.engine.device, dtype=dtype).contiguous()
        return t.from_numpy(np.ascontiguousarray(np.asarray(a))).to(device=self.engine.device, dtype=dtype)

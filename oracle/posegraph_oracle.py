"""CPU oracle for the camera-object pose-graph edges.  TEST INFRASTRUCTURE:
imported only by tests/, never by the product path.

PARITY UNPINNED.  The edge class is the reference's own (`include/ObjectPoseGraph.h:56-88`), but everything it calls
is third-party C++ that cannot be built in this image (no Eigen, no cmake build of g2o):
  * g2o, vendored by the reference under `Thirdparty/g2o` (SE3Quat: `g2o/types/se3quat.h:56-286`, `deltaR`/`skew`:
    `g2o/types/se3_ops.hpp`, Huber kernel: `g2o/core/robust_kernel_impl.cpp:78-91`, quadratic form:
    `g2o/core/base_binary_edge.hpp:60-115`, `robustInformation`: `g2o/core/base_edge.h:96-102`);
  * Eigen 3.4.0 (`CMakeLists.txt:38`, `build_cuda102.sh:119`), absent from /root/reference: quaternion from a rotation
    matrix, quaternion product, rotation of a vector, `toRotationMatrix`, `normalize` are restated from Eigen's
    published Geometry module (`Quaternion.h`: `quaternionbase_assign_impl<Other,3,3>`, `quat_product`,
    `_transformVector`, `toRotationMatrix`).
With no reference binary and no golden vectors for this path (the reference has no tests), the restatement is checked
only through properties (tests/test_pose_graph.py): log(exp(x)) = x, quaternion <-> matrix round trips, the adjoint
identity, the analytic Jacobians against central differences under g2o's `oplus`, the Huber weights at the threshold.

Everything is FP64 scalar arithmetic in the order the C++ evaluates it.  Poses are g2o's `Vector7d` layout:
(tx, ty, tz, qx, qy, qz, qw).
"""
from __future__ import annotations

import math

import numpy as np


# ---------------------------------------------------------------------------------------------------------------
# Eigen 3.4.0 Quaternion pieces
# ---------------------------------------------------------------------------------------------------------------
def quat_from_matrix(R):
    """Eigen `Quaterniond(Matrix3d)`; returns (x, y, z, w)."""
    t = R[0][0] + R[1][1] + R[2][2]
    q = [0.0, 0.0, 0.0, 0.0]
    if t > 0.0:
        t = math.sqrt(t + 1.0)
        q[3] = 0.5 * t
        t = 0.5 / t
        q[0] = (R[2][1] - R[1][2]) * t
        q[1] = (R[0][2] - R[2][0]) * t
        q[2] = (R[1][0] - R[0][1]) * t
    else:
        i = 0
        if R[1][1] > R[0][0]:
            i = 1
        if R[2][2] > R[i][i]:
            i = 2
        j = (i + 1) % 3
        k = (j + 1) % 3
        t = math.sqrt(R[i][i] - R[j][j] - R[k][k] + 1.0)
        q[i] = 0.5 * t
        t = 0.5 / t
        q[3] = (R[k][j] - R[j][k]) * t
        q[j] = (R[j][i] + R[i][j]) * t
        q[k] = (R[k][i] + R[i][k]) * t
    return q


def quat_mul(a, b):
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return [aw * bx + ax * bw + ay * bz - az * by,
            aw * by + ay * bw + az * bx - ax * bz,
            aw * bz + az * bw + ax * by - ay * bx,
            aw * bw - ax * bx - ay * by - az * bz]


def quat_rotate(q, v):
    """Eigen `_transformVector`: uv = 2 (q.vec x v); v + w uv + q.vec x uv."""
    x, y, z, w = q
    uv = [y * v[2] - z * v[1], z * v[0] - x * v[2], x * v[1] - y * v[0]]
    uv = [uv[0] + uv[0], uv[1] + uv[1], uv[2] + uv[2]]
    c = [y * uv[2] - z * uv[1], z * uv[0] - x * uv[2], x * uv[1] - y * uv[0]]
    return [v[0] + w * uv[0] + c[0], v[1] + w * uv[1] + c[1], v[2] + w * uv[2] + c[2]]


def quat_to_matrix(q):
    x, y, z, w = q
    tx, ty, tz = 2.0 * x, 2.0 * y, 2.0 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return [[1.0 - (tyy + tzz), txy - twz, txz + twy],
            [txy + twz, 1.0 - (txx + tzz), tyz - twx],
            [txz - twy, tyz + twx, 1.0 - (txx + tyy)]]


def quat_normalized(q):
    """SE3Quat::normalizeRotation (se3quat.h:280-285): w >= 0, unit norm."""
    if q[3] < 0.0:
        q = [-c for c in q]
    n = math.sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3])
    return [c / n for c in q]


# ---------------------------------------------------------------------------------------------------------------
# g2o SE3Quat
# ---------------------------------------------------------------------------------------------------------------
def se3_from_matrix(T):
    """Converter::toSE3Quat (src/Converter.cc:37-54): FP32 4x4 -> double R, t -> SE3Quat(R, t)."""
    T = np.asarray(T, np.float32).astype(np.float64)
    q = quat_normalized(quat_from_matrix(T[:3, :3].tolist()))
    return [float(T[0, 3]), float(T[1, 3]), float(T[2, 3])] + q


def se3_mul(a, b):
    """se3quat.h:104-110."""
    r = quat_rotate(a[3:], b[:3])
    t = [a[0] + r[0], a[1] + r[1], a[2] + r[2]]
    return t + quat_normalized(quat_mul(a[3:], b[3:]))


def se3_inv(a):
    """se3quat.h:123-128 (no re-normalisation)."""
    qc = [-a[3], -a[4], -a[5], a[6]]
    t = quat_rotate(qc, [a[0] * -1.0, a[1] * -1.0, a[2] * -1.0])
    return t + qc


def _skew(v):
    return [[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]]


def _mm(A, B):
    return [[A[i][0] * B[0][j] + A[i][1] * B[1][j] + A[i][2] * B[2][j] for j in range(3)] for i in range(3)]


def se3_log(a):
    """se3quat.h:178-215 -> (omega, upsilon)."""
    R = quat_to_matrix(a[3:])
    d = 0.5 * (R[0][0] + R[1][1] + R[2][2] - 1.0)
    dR = [R[2][1] - R[1][2], R[0][2] - R[2][0], R[1][0] - R[0][1]]
    if d > 0.99999:
        om = [0.5 * c for c in dR]
        c2 = 1.0 / 12.0
    else:
        theta = math.acos(d)
        f = theta / (2.0 * math.sqrt(1.0 - d * d))
        om = [f * c for c in dR]
        c2 = (1.0 - theta / (2.0 * math.tan(theta / 2.0))) / (theta * theta)
    O = _skew(om)
    O2 = _mm(O, O)
    Vi = [[(1.0 if i == j else 0.0) - 0.5 * O[i][j] + c2 * O2[i][j] for j in range(3)] for i in range(3)]
    up = [Vi[i][0] * a[0] + Vi[i][1] * a[1] + Vi[i][2] * a[2] for i in range(3)]
    return om + up


def se3_exp(u):
    """se3quat.h:223-257, update = (omega, upsilon)."""
    om, up = list(u[:3]), list(u[3:6])
    theta = math.sqrt(om[0] * om[0] + om[1] * om[1] + om[2] * om[2])
    O = _skew(om)
    O2 = _mm(O, O)
    eye = [[1.0 if i == j else 0.0 for j in range(3)] for i in range(3)]
    if theta < 0.00001:
        R = [[eye[i][j] + O[i][j] + O2[i][j] for j in range(3)] for i in range(3)]
        V = R
    else:
        a = math.sin(theta) / theta
        b = (1.0 - math.cos(theta)) / (theta * theta)
        c = (theta - math.sin(theta)) / (theta ** 3)
        R = [[eye[i][j] + a * O[i][j] + b * O2[i][j] for j in range(3)] for i in range(3)]
        V = [[eye[i][j] + b * O[i][j] + c * O2[i][j] for j in range(3)] for i in range(3)]
    t = [V[i][0] * up[0] + V[i][1] * up[1] + V[i][2] * up[2] for i in range(3)]
    return t + quat_normalized(quat_from_matrix(R))


def se3_adj(a):
    """se3quat.h:259-268: [[R, 0], [skew(t) R, R]]."""
    R = quat_to_matrix(a[3:])
    tR = _mm(_skew(a[:3]), R)
    A = np.zeros((6, 6))
    A[:3, :3] = R
    A[3:, 3:] = R
    A[3:, :3] = tR
    return A


def se3_oplus(a, u):
    """VertexSE3Expmap::oplusImpl (g2o/types/types_six_dof_expmap.h): estimate <- exp(update) * estimate."""
    return se3_mul(se3_exp(u), a)


# ---------------------------------------------------------------------------------------------------------------
# the edge (include/ObjectPoseGraph.h:56-88) and its quadratic form (g2o base_binary_edge.hpp:60-115)
# ---------------------------------------------------------------------------------------------------------------
def edge_error(meas, vi, vj):
    """computeError: log(Tij^-1 * Ti * Tj^-1)."""
    return se3_log(se3_mul(se3_mul(se3_inv(meas), vi), se3_inv(vj)))


def edge_jacobians(meas, err):
    """linearizeOplus: J = 0.5 [[skew(w), 0], [skew(t), skew(w)]] + I;  Ji = J adj(Tij^-1), Jj = -J."""
    w, t = err[:3], err[3:]
    J = np.zeros((6, 6))
    J[:3, :3] = _skew(w)
    J[3:, :3] = _skew(t)
    J[3:, 3:] = _skew(w)
    J = 0.5 * J + np.eye(6)
    return J @ se3_adj(se3_inv(meas)), -J


def huber(chi2, delta):
    """RobustKernelHuber::robustify -> (rho, rho')."""
    if delta <= 0.0:
        return chi2, 1.0
    dsqr = delta * delta
    if chi2 <= dsqr:
        return chi2, 1.0
    s = math.sqrt(chi2)
    return 2.0 * s * delta - dsqr, delta / s


def linearize(verts, edge_i, edge_j, meas, inv_sigma, huber_delta):
    """All edges: error, Ji, Jj, chi2 (= e^T Omega e, Omega = inv_sigma I), rho, rho', and the quadratic form
    accumulated per vertex in edge order: A[v] += J^T (rho' Omega) J, b[v] += J^T (-rho' Omega e), plus the
    per-edge off-diagonal block Ji^T (rho' Omega) Jj."""
    nv, ne = len(verts), len(meas)
    out = dict(err=np.zeros((ne, 6)), Ji=np.zeros((ne, 6, 6)), Jj=np.zeros((ne, 6, 6)), chi2=np.zeros(ne),
               rho=np.zeros(ne), w=np.zeros(ne), A=np.zeros((nv, 6, 6)), b=np.zeros((nv, 6)), Hij=np.zeros((ne, 6, 6)))
    for e in range(ne):
        i, j = int(edge_i[e]), int(edge_j[e])
        err = edge_error(list(meas[e]), list(verts[i]), list(verts[j]))
        Ji, Jj = edge_jacobians(list(meas[e]), err)
        err = np.array(err)
        chi2 = float(err @ (inv_sigma * err))
        rho, w = huber(chi2, huber_delta)
        om = w * inv_sigma
        r = -om * err
        out["err"][e], out["Ji"][e], out["Jj"][e] = err, Ji, Jj
        out["chi2"][e], out["rho"][e], out["w"][e] = chi2, rho, w
        out["A"][i] += Ji.T @ (om * Ji)
        out["A"][j] += Jj.T @ (om * Jj)
        out["b"][i] += Ji.T @ r
        out["b"][j] += Jj.T @ r
        out["Hij"][e] = Ji.T @ (om * Jj)
    return out

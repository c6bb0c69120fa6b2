// Camera-object pose-graph edges of the joint bundle adjustment: the consumer of
// the SE(3) measurement that dspslam_gn_pose_only / dspslam_gn_joint produce.
//
//   reference: include/ObjectPoseGraph.h:56-88 (EdgeSE3LieAlgebra::computeError / linearizeOplus), built in
//   src/Optimizer_util.cc:199-225 and :555-583 (information = invSigmaObject * I, Huber kernel), read back through
//   chi2() at :647-656 and :701-711.  The Lie-group arithmetic underneath is g2o's SE3Quat
//   (Thirdparty/g2o/g2o/types/se3quat.h) on Eigen 3.4.0 quaternions; both restated here in FP64, operation by
//   operation (see oracle/posegraph_oracle.py for the citations of every piece).
//
// Everything is FP64 and tiny: one thread per edge for error + Jacobians + robust weight, one small CTA per vertex
// for the quadratic form (each thread owns one entry of the 6x6 block or of the 6-vector and walks the edges in
// order, so the sums are order-stable).  Latency-bound by construction; the point of having it on the device is
// that the measurement never leaves it.
#include <cuda_runtime.h>
#include <stdint.h>

#include "posegraph.h"

namespace dsp {
namespace {

struct Q { double x, y, z, w; };
struct P7 { double t[3]; Q q; };

__device__ __forceinline__ P7 load7(const double* p) {
    P7 a;
    a.t[0] = p[0]; a.t[1] = p[1]; a.t[2] = p[2];
    a.q.x = p[3]; a.q.y = p[4]; a.q.z = p[5]; a.q.w = p[6];
    return a;
}
__device__ __forceinline__ void store7(double* p, const P7& a) {
    p[0] = a.t[0]; p[1] = a.t[1]; p[2] = a.t[2];
    p[3] = a.q.x; p[4] = a.q.y; p[5] = a.q.z; p[6] = a.q.w;
}

// Eigen Quaterniond(Matrix3d)
__device__ Q q_from_R(const double R[3][3]) {
    double t = R[0][0] + R[1][1] + R[2][2];
    double q[4];
    if (t > 0.0) {
        t = sqrt(t + 1.0);
        q[3] = 0.5 * t;
        t = 0.5 / t;
        q[0] = (R[2][1] - R[1][2]) * t;
        q[1] = (R[0][2] - R[2][0]) * t;
        q[2] = (R[1][0] - R[0][1]) * t;
    } else {
        int i = 0;
        if (R[1][1] > R[0][0]) i = 1;
        if (R[2][2] > R[i][i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = sqrt(R[i][i] - R[j][j] - R[k][k] + 1.0);
        q[i] = 0.5 * t;
        t = 0.5 / t;
        q[3] = (R[k][j] - R[j][k]) * t;
        q[j] = (R[j][i] + R[i][j]) * t;
        q[k] = (R[k][i] + R[i][k]) * t;
    }
    return Q{q[0], q[1], q[2], q[3]};
}

__device__ __forceinline__ Q q_mul(const Q& a, const Q& b) {
    return Q{a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
             a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}

// Eigen _transformVector
__device__ __forceinline__ void q_rot(const Q& q, const double v[3], double o[3]) {
    double u0 = q.y * v[2] - q.z * v[1], u1 = q.z * v[0] - q.x * v[2], u2 = q.x * v[1] - q.y * v[0];
    u0 += u0; u1 += u1; u2 += u2;
    o[0] = v[0] + q.w * u0 + (q.y * u2 - q.z * u1);
    o[1] = v[1] + q.w * u1 + (q.z * u0 - q.x * u2);
    o[2] = v[2] + q.w * u2 + (q.x * u1 - q.y * u0);
}

__device__ __forceinline__ void q_to_R(const Q& q, double R[3][3]) {
    const double tx = 2.0 * q.x, ty = 2.0 * q.y, tz = 2.0 * q.z;
    const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
    const double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    R[0][0] = 1.0 - (tyy + tzz); R[0][1] = txy - twz;         R[0][2] = txz + twy;
    R[1][0] = txy + twz;         R[1][1] = 1.0 - (txx + tzz); R[1][2] = tyz - twx;
    R[2][0] = txz - twy;         R[2][1] = tyz + twx;         R[2][2] = 1.0 - (txx + tyy);
}

// SE3Quat::normalizeRotation
__device__ __forceinline__ Q q_norm(Q q) {
    if (q.w < 0.0) { q.x = -q.x; q.y = -q.y; q.z = -q.z; q.w = -q.w; }
    const double n = sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    return Q{q.x / n, q.y / n, q.z / n, q.w / n};
}

__device__ P7 se3_mul(const P7& a, const P7& b) {
    P7 r;
    double o[3];
    q_rot(a.q, b.t, o);
    r.t[0] = a.t[0] + o[0]; r.t[1] = a.t[1] + o[1]; r.t[2] = a.t[2] + o[2];
    r.q = q_norm(q_mul(a.q, b.q));
    return r;
}

__device__ P7 se3_inv(const P7& a) {
    P7 r;
    r.q = Q{-a.q.x, -a.q.y, -a.q.z, a.q.w};
    const double m[3] = {a.t[0] * -1.0, a.t[1] * -1.0, a.t[2] * -1.0};
    q_rot(r.q, m, r.t);
    return r;
}

__device__ __forceinline__ void skew(const double v[3], double S[3][3]) {
    S[0][0] = 0.0;   S[0][1] = -v[2]; S[0][2] = v[1];
    S[1][0] = v[2];  S[1][1] = 0.0;   S[1][2] = -v[0];
    S[2][0] = -v[1]; S[2][1] = v[0];  S[2][2] = 0.0;
}
__device__ __forceinline__ void mm3(const double A[3][3], const double B[3][3], double C[3][3]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) C[i][j] = A[i][0] * B[0][j] + A[i][1] * B[1][j] + A[i][2] * B[2][j];
}

// SE3Quat::log -> (omega, upsilon)
__device__ void se3_log(const P7& a, double e[6]) {
    double R[3][3];
    q_to_R(a.q, R);
    const double d = 0.5 * (R[0][0] + R[1][1] + R[2][2] - 1.0);
    const double dR[3] = {R[2][1] - R[1][2], R[0][2] - R[2][0], R[1][0] - R[0][1]};
    double om[3], c2;
    if (d > 0.99999) {
        for (int i = 0; i < 3; ++i) om[i] = 0.5 * dR[i];
        c2 = 1.0 / 12.0;
    } else {
        const double theta = acos(d);
        const double f = theta / (2.0 * sqrt(1.0 - d * d));
        for (int i = 0; i < 3; ++i) om[i] = f * dR[i];
        c2 = (1.0 - theta / (2.0 * tan(theta / 2.0))) / (theta * theta);
    }
    double O[3][3], O2[3][3];
    skew(om, O);
    mm3(O, O, O2);
    for (int i = 0; i < 3; ++i) {
        double vi[3];
        for (int j = 0; j < 3; ++j) vi[j] = (i == j ? 1.0 : 0.0) - 0.5 * O[i][j] + c2 * O2[i][j];
        e[i] = om[i];
        e[3 + i] = vi[0] * a.t[0] + vi[1] * a.t[1] + vi[2] * a.t[2];
    }
}

// SE3Quat::exp, update = (omega, upsilon)
__device__ P7 se3_exp(const double u[6]) {
    const double om[3] = {u[0], u[1], u[2]};
    const double theta = sqrt(om[0] * om[0] + om[1] * om[1] + om[2] * om[2]);
    double O[3][3], O2[3][3], R[3][3], V[3][3];
    skew(om, O);
    mm3(O, O, O2);
    if (theta < 0.00001) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) V[i][j] = R[i][j] = (i == j ? 1.0 : 0.0) + O[i][j] + O2[i][j];
    } else {
        const double a = sin(theta) / theta, b = (1.0 - cos(theta)) / (theta * theta),
                     c = (theta - sin(theta)) / (theta * theta * theta);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) {
                const double id = (i == j ? 1.0 : 0.0);
                R[i][j] = id + a * O[i][j] + b * O2[i][j];
                V[i][j] = id + b * O[i][j] + c * O2[i][j];
            }
    }
    P7 r;
    for (int i = 0; i < 3; ++i) r.t[i] = V[i][0] * u[3] + V[i][1] * u[4] + V[i][2] * u[5];
    r.q = q_norm(q_from_R(R));
    return r;
}

// Converter::toSE3Quat: FP32 4x4 (row-major) -> SE3Quat(R, t)
__global__ void k_pg_from_matrices(const float* __restrict__ T, int n, double* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const float* m = T + (size_t)e * 16;
    double R[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[i][j] = (double)m[i * 4 + j];
    P7 a;
    a.t[0] = (double)m[3]; a.t[1] = (double)m[7]; a.t[2] = (double)m[11];
    a.q = q_norm(q_from_R(R));
    store7(out + (size_t)e * 7, a);
}

// computeError + linearizeOplus + chi2 + Huber, one thread per edge
__global__ void k_pg_edges(const double* __restrict__ verts, const int32_t* __restrict__ ei,
                           const int32_t* __restrict__ ej, const double* __restrict__ meas, int ne, double inv_sigma,
                           double huber_delta, double* __restrict__ err, double* __restrict__ Ji,
                           double* __restrict__ Jj, double* __restrict__ chi2, double* __restrict__ rho,
                           double* __restrict__ wgt) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ne) return;
    const P7 Mi = se3_inv(load7(meas + (size_t)e * 7));
    const P7 E = se3_mul(se3_mul(Mi, load7(verts + (size_t)ei[e] * 7)), se3_inv(load7(verts + (size_t)ej[e] * 7)));
    double er[6];
    se3_log(E, er);
    // J = 0.5 [[skew(w), 0], [skew(t), skew(w)]] + I
    double Sw[3][3], St[3][3], J[6][6];
    skew(er, Sw);
    skew(er + 3, St);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const double id = (i == j ? 1.0 : 0.0);
            J[i][j] = 0.5 * Sw[i][j] + id;
            J[i][3 + j] = 0.0;
            J[3 + i][j] = 0.5 * St[i][j];
            J[3 + i][3 + j] = 0.5 * Sw[i][j] + id;
        }
    // adj(Tij^-1) = [[R, 0], [skew(t) R, R]]
    double R[3][3], Tx[3][3], TR[3][3];
    q_to_R(Mi.q, R);
    skew(Mi.t, Tx);
    mm3(Tx, R, TR);
    double* ji = Ji + (size_t)e * 36;
    double* jj = Jj + (size_t)e * 36;
    for (int i = 0; i < 6; ++i)
        for (int j = 0; j < 6; ++j) {
            double s = 0.0;
            for (int k = 0; k < 6; ++k) {
                const double a = (k < 3) ? (j < 3 ? R[k][j] : 0.0) : (j < 3 ? TR[k - 3][j] : R[k - 3][j - 3]);
                s += J[i][k] * a;
            }
            ji[i * 6 + j] = s;
            jj[i * 6 + j] = -J[i][j];
        }
    double c = 0.0;
    for (int i = 0; i < 6; ++i) {
        err[(size_t)e * 6 + i] = er[i];
        c += er[i] * (inv_sigma * er[i]);
    }
    double r = c, w = 1.0;
    if (huber_delta > 0.0 && c > huber_delta * huber_delta) {           // RobustKernelHuber::robustify
        const double s = sqrt(c);
        r = 2.0 * s * huber_delta - huber_delta * huber_delta;
        w = huber_delta / s;
    }
    chi2[e] = c;
    rho[e] = r;
    wgt[e] = w;
}

// constructQuadraticForm, gathered per vertex: thread q < 36 owns A[q/6][q%6], threads 36..41 own b.
// The edge list is walked in chunks of PG_CHUNK: each of the two warps compacts the incident edges of its half of
// the chunk (ballot + prefix count, order kept) into shared memory, then the 42 owners add them up - first half,
// then second half, i.e. in edge order, so the sums do not depend on the launch shape.
constexpr int PG_CHUNK = 2048;

__global__ void __launch_bounds__(64) k_pg_vertex_blocks(const int32_t* __restrict__ ei, const int32_t* __restrict__ ej,
                                                         int ne, double inv_sigma, const double* __restrict__ err,
                                                         const double* __restrict__ Ji, const double* __restrict__ Jj,
                                                         const double* __restrict__ wgt, double* __restrict__ A,
                                                         double* __restrict__ b) {
    __shared__ int32_t s_list[2][PG_CHUNK / 2];      // edge id, bit 31 set when the vertex is the edge's j end
    __shared__ int s_cnt[2];
    const int v = blockIdx.x, q = threadIdx.x, warp = q >> 5, lane = q & 31;
    double acc = 0.0;
    for (int c0 = 0; c0 < ne; c0 += PG_CHUNK) {
        const int lo = c0 + warp * (PG_CHUNK / 2), hi = min(lo + PG_CHUNK / 2, ne);
        int cnt = 0;
        for (int e0 = lo; e0 < hi; e0 += 32) {
            const int e = e0 + lane;
            const bool is_i = e < hi && ei[e] == v, is_j = e < hi && ej[e] == v;
            const unsigned m = __ballot_sync(0xffffffffu, is_i || is_j);
            if (is_i || is_j) s_list[warp][cnt + __popc(m & ((1u << lane) - 1))] = is_i ? e : (e | (int32_t)0x80000000);
            cnt += __popc(m);
        }
        if (lane == 0) s_cnt[warp] = cnt;
        __syncthreads();
        if (q < 42) {
            for (int h = 0; h < 2; ++h)
                for (int k = 0; k < s_cnt[h]; ++k) {
                    const int32_t code = s_list[h][k];
                    const int e = code & 0x7fffffff;
                    const double* J = (code < 0 ? Jj : Ji) + (size_t)e * 36;
                    const double om = wgt[e] * inv_sigma;
                    double s = 0.0;
                    if (q < 36) {
                        const int r = q / 6, c = q % 6;
                        for (int t = 0; t < 6; ++t) s += J[t * 6 + r] * (om * J[t * 6 + c]);
                    } else {
                        const int r = q - 36;
                        for (int t = 0; t < 6; ++t) s += J[t * 6 + r] * (-om * err[(size_t)e * 6 + t]);
                    }
                    acc += s;
                }
        }
        __syncthreads();
    }
    if (q < 36) A[(size_t)v * 36 + q] = acc; else if (q < 42) b[(size_t)v * 6 + q - 36] = acc;
}

// off-diagonal block of every edge: Ji^T (rho' Omega) Jj
__global__ void k_pg_edge_blocks(int ne, double inv_sigma, const double* __restrict__ Ji,
                                 const double* __restrict__ Jj, const double* __restrict__ wgt,
                                 double* __restrict__ Hij) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ne * 36) return;
    const int e = t / 36, r = (t % 36) / 6, c = t % 6;
    const double om = wgt[e] * inv_sigma;
    const double* a = Ji + (size_t)e * 36;
    const double* bb = Jj + (size_t)e * 36;
    double s = 0.0;
    for (int k = 0; k < 6; ++k) s += a[k * 6 + r] * (om * bb[k * 6 + c]);
    Hij[t] = s;
}

// VertexSE3Expmap::oplusImpl: estimate <- exp(update) * estimate
__global__ void k_pg_oplus(double* __restrict__ verts, const double* __restrict__ upd, int n) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    double u[6];
    for (int i = 0; i < 6; ++i) u[i] = upd[(size_t)v * 6 + i];
    store7(verts + (size_t)v * 7, se3_mul(se3_exp(u), load7(verts + (size_t)v * 7)));
}

}  // namespace

int launch_pg_from_matrices(const float* T, int n, double* out, cudaStream_t st) {
    if (n == 0) return 0;
    k_pg_from_matrices<<<(n + 127) / 128, 128, 0, st>>>(T, n, out);
    return 1;
}

int launch_pg_linearize(const double* verts, int nv, const int32_t* ei, const int32_t* ej, const double* meas, int ne,
                        double inv_sigma, double huber_delta, double* err, double* Ji, double* Jj, double* chi2,
                        double* rho, double* wgt, double* A, double* b, double* Hij, cudaStream_t st) {
    int nl = 0;
    if (ne > 0) {
        k_pg_edges<<<(ne + 63) / 64, 64, 0, st>>>(verts, ei, ej, meas, ne, inv_sigma, huber_delta, err, Ji, Jj, chi2,
                                                  rho, wgt);
        ++nl;
        if (Hij) {
            k_pg_edge_blocks<<<(ne * 36 + 255) / 256, 256, 0, st>>>(ne, inv_sigma, Ji, Jj, wgt, Hij);
            ++nl;
        }
    }
    if (A && b && nv > 0) {
        k_pg_vertex_blocks<<<nv, 64, 0, st>>>(ei, ej, ne, inv_sigma, err, Ji, Jj, wgt, A, b);
        ++nl;
    }
    return nl;
}

int launch_pg_oplus(double* verts, const double* upd, int n, cudaStream_t st) {
    if (n == 0) return 0;
    k_pg_oplus<<<(n + 127) / 128, 128, 0, st>>>(verts, upd, n);
    return 1;
}

}  // namespace dsp

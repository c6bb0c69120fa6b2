// Host/device scalar math of the Gauss-Newton path: Lie-group exponentials, the
// Sim(3) inverse, the rotation prior, the Huber re-weighting and the per-ray
// render scan.  Everything here is `__host__ __device__` so that the exact code
// the kernels execute can also be compiled with g++ and checked against the
// oracle on a CPU-only box (tests/test_hd_math.py builds csrc/hd_selftest.cpp).
//
// Reference lines restated:
//   exp_se3   reconstruct/loss_utils.py:139-173
//   exp_sim3  reconstruct/loss_utils.py:198-243 (incl. the c=0 quirk for sigma<=eps)
//   huber     reconstruct/loss_utils.py:246-275
//   rot prior reconstruct/loss.py:169-192
//   ray scan  reconstruct/loss.py:98-155
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define DSP_HD __host__ __device__ __forceinline__
#else
#define DSP_HD inline
#endif

namespace dsp {

// ---- 3x3 / 4x4 helpers (row-major) ----------------------------------------
DSP_HD float det3(const float* a) {
    return a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) +
           a[2] * (a[3] * a[7] - a[4] * a[6]);
}

// inverse of an affine 4x4 [A t; 0 1] (row-major 16 floats); A general 3x3.
// Computed in double and rounded once (the reference uses FP32 LU; both are
// within FP32 rounding of the exact inverse).
DSP_HD void inv_affine(const float* T, float* Ti) {
    double a[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) a[i * 3 + j] = (double)T[i * 4 + j];
    double c00 = a[4] * a[8] - a[5] * a[7], c01 = a[5] * a[6] - a[3] * a[8], c02 = a[3] * a[7] - a[4] * a[6];
    double det = a[0] * c00 + a[1] * c01 + a[2] * c02;
    double id = 1.0 / det;
    double inv[9];
    inv[0] = c00 * id;
    inv[1] = (a[2] * a[7] - a[1] * a[8]) * id;
    inv[2] = (a[1] * a[5] - a[2] * a[4]) * id;
    inv[3] = c01 * id;
    inv[4] = (a[0] * a[8] - a[2] * a[6]) * id;
    inv[5] = (a[2] * a[3] - a[0] * a[5]) * id;
    inv[6] = c02 * id;
    inv[7] = (a[1] * a[6] - a[0] * a[7]) * id;
    inv[8] = (a[0] * a[4] - a[1] * a[3]) * id;
    double t0 = T[3], t1 = T[7], t2 = T[11];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) Ti[i * 4 + j] = (float)inv[i * 3 + j];
        Ti[i * 4 + 3] = (float)(-(inv[i * 3 + 0] * t0 + inv[i * 3 + 1] * t1 + inv[i * 3 + 2] * t2));
    }
    Ti[12] = 0.f; Ti[13] = 0.f; Ti[14] = 0.f; Ti[15] = 1.f;
}

// C = A * B for affine 4x4 (row-major, last row 0 0 0 1)
DSP_HD void mul_affine(const float* A, const float* B, float* C) {
    float r[16];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 4; ++j) {
            float s = A[i * 4 + 0] * B[0 * 4 + j] + A[i * 4 + 1] * B[1 * 4 + j] + A[i * 4 + 2] * B[2 * 4 + j];
            if (j == 3) s += A[i * 4 + 3];
            r[i * 4 + j] = s;
        }
    }
    r[12] = 0.f; r[13] = 0.f; r[14] = 0.f; r[15] = 1.f;
    for (int i = 0; i < 16; ++i) C[i] = r[i];
}

DSP_HD void hat_and_square(const float* w, float* wh, float* wh2) {
    wh[0] = 0.f;   wh[1] = -w[2]; wh[2] = w[1];
    wh[3] = w[2];  wh[4] = 0.f;   wh[5] = -w[0];
    wh[6] = -w[1]; wh[7] = w[0];  wh[8] = 0.f;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            wh2[i * 3 + j] = wh[i * 3 + 0] * wh[0 * 3 + j] + wh[i * 3 + 1] * wh[1 * 3 + j] + wh[i * 3 + 2] * wh[2 * 3 + j];
}

// exp_se3: x = (v, w) -> 4x4.   loss_utils.py:139-173
DSP_HD void exp_se3(const float* x, float* T) {
    const float* v = x;
    const float* w = x + 3;
    float wh[9], wh2[9], ew[9], J[9];
    hat_and_square(w, wh, wh2);
    float theta = sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    if (theta <= 1e-8f) {
        for (int i = 0; i < 9; ++i) { ew[i] = (i % 4 == 0) ? 1.f : 0.f; J[i] = ew[i]; }
    } else {
        float st = sinf(theta), ct = cosf(theta);
        float th2 = theta * theta, th3 = th2 * theta;
        float k1 = (1.f - ct) / th2, k2 = (theta - st) / th3;
        for (int i = 0; i < 9; ++i) {
            float e = (i % 4 == 0) ? 1.f : 0.f;
            ew[i] = e + wh[i] * st / theta + wh2[i] * (1.f - ct) / th2;
            J[i] = e + k1 * wh[i] + k2 * wh2[i];
        }
    }
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) T[i * 4 + j] = ew[i * 3 + j];
        T[i * 4 + 3] = J[i * 3 + 0] * v[0] + J[i * 3 + 1] * v[1] + J[i * 3 + 2] * v[2];
    }
    T[12] = 0.f; T[13] = 0.f; T[14] = 0.f; T[15] = 1.f;
}

// exp_sim3: x = (v, w, sigma) -> 4x4.   loss_utils.py:198-243
DSP_HD void exp_sim3(const float* x, float* T) {
    const float* v = x;
    const float* w = x + 3;
    float s = x[6];
    float wh[9], wh2[9], ew[9], J[9];
    hat_and_square(w, wh, wh2);
    float theta = sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    float th2 = theta * theta;
    float st = sinf(theta), ct = cosf(theta);
    float es = expf(s);
    float s2 = s * s;
    if (theta <= 1e-8f) {
        float c = (s == 0.f) ? 1.f : (es - 1.f) / s;
        for (int i = 0; i < 9; ++i) { ew[i] = (i % 4 == 0) ? 1.f : 0.f; J[i] = c * ew[i]; }
    } else {
        float a = es * st, b = es * ct;
        float c = (s <= 1e-8f) ? 0.f : (es - 1.f) / s;          // reference quirk (:233)
        float k1 = (a * s + (1.f - b) * theta) / (s2 + th2);
        float k2 = c - ((b - 1.f) * s + a * theta) / (s2 + th2);
        for (int i = 0; i < 9; ++i) {
            float e = (i % 4 == 0) ? 1.f : 0.f;
            ew[i] = e + wh[i] * st / theta + wh2[i] * (1.f - ct) / th2;
            J[i] = c * e + k1 * wh[i] / theta + k2 * wh2[i] / th2;
        }
    }
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) T[i * 4 + j] = es * ew[i * 3 + j];
        T[i * 4 + 3] = J[i * 3 + 0] * v[0] + J[i * 3 + 1] * v[1] + J[i * 3 + 2] * v[2];
    }
    T[12] = 0.f; T[13] = 0.f; T[14] = 0.f; T[15] = 1.f;
}

// Huber-scaled residual: returns w*r with w = 1 (|r|<=b), sqrt(2b|r|-b^2)/|r| otherwise,
// w = 0 at r == 0.   loss_utils.py:246-275
DSP_HD float huber_res(float r, float b) {
    float x = fabsf(r);
    float rn = (x <= b) ? x * x : 2.f * b * x - b * b;
    float den = (x == 0.f) ? 1.f : x;
    return sqrtf(rn) / den * r;
}

// Rotation (upright) prior.  loss.py:169-192.  T_co = inverse(T_oc) is passed in.
// Writes J[3] (entries 3..5 of the 7-vector) and returns the residual (0 => inactive).
DSP_HD float rot_prior(const float* T_co, float scale, float* J3) {
    float r[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) r[i * 3 + j] = T_co[i * 4 + j] / scale;
    // ry = R_co * e_y = second column; n_g = (0,-1,0)  ->  res = 1 + r[1][1]
    float res = 1.f - (-r[4]);
    J3[0] = J3[1] = J3[2] = 0.f;
    if (res < 1e-7f) return 0.f;
    // R_oc = inverse(R_co) (general inverse, as the reference does), a = R_oc * n_g = -(column 1 of R_oc)
    float c10 = r[2] * r[7] - r[1] * r[8];
    float c11 = r[0] * r[8] - r[2] * r[6];
    float c12 = r[1] * r[6] - r[0] * r[7];
    float d = det3(r);
    float a0 = -c10 / d, a1 = -c11 / d, a2 = -c12 / d;
    (void)a1;
    // J = a x e_y = (a1*0 - a2*1, a2*0 - a0*0, a0*1 - a1*0) = (-a2, 0, a0)
    J3[0] = -a2;
    J3[1] = 0.f;
    J3[2] = a0;
    return res;
}

// torch.linspace element i of n between a and b: symmetric evaluation, carried out in double
// and rounded once (what torch's CPU kernel yields for a float32 result).
DSP_HD float linspace_at(float a, float b, int n, int i) {
    const double step = ((double)b - (double)a) / (double)(n - 1);
    return (float)((i < n / 2) ? (double)a + step * (double)i : (double)b - step * (double)(n - 1 - i));
}

// ---- per-ray render scan ----------------------------------------------------
// loss.py:98-155 for ONE ray.  `sdf[j]` holds the decoder output of sample j, or a
// huge positive sentinel for samples outside the unit ball (occupancy 0, no grad).
// Selected samples (|s|<th and de_do>1e-2) are reported through `emit(j, de_ds, res)`
// in increasing j.  Returns the number selected.
struct RayParams {
    float th;          // cut-off threshold
    float dmin, dmax;  // depth range (sampled_depth[0], sampled_depth[D-1])
    int D;
    const float* dtab; // null, or the D sampled depths (linspace_at(dmin, dmax, D, j)) precomputed per object
};

#define DSP_MAX_D 64

DSP_HD float ray_depth(const RayParams& p, int j) {
    return p.dtab ? p.dtab[j] : linspace_at(p.dmin, p.dmax, p.D, j);
}

// `acc` is scratch for D floats addressed as acc[j * acc_stride] (the CUDA kernel hands in a shared-memory column:
// a dynamically indexed local array would live in local memory).
template <class Emit>
DSP_HD int ray_scan_scratch(const float* sdf, int stride, float depth_obs, const RayParams& p, float* acc_base,
                            int acc_stride, Emit emit) {
    const int D = p.D;
    struct Acc {
        float* b; int s;
        DSP_HD float& operator[](int j) const { return b[j * s]; }
    } acc{acc_base, acc_stride};
    // pass 1: transmittance acc_j = prod_{i<=j} (1-o_i) and expected depth
    //         d_u = sum_j d_j * o_j * acc_{j-1} + 1.1*dmax * acc_{D-1}
    // Once the transmittance is exactly 0 every later term is exactly 0: stop there (jend).
    float T = 1.f, d_u = 0.f;
    int jend = D;
    for (int j = 0; j < D; ++j) {
        float s = sdf[j * stride];
        float o = 0.5f - fminf(fmaxf(s, -p.th), p.th) / (2.f * p.th);
        d_u += ray_depth(p, j) * (o * T);
        T *= (1.f - o);
        acc[j] = T;
        if (T == 0.f) { jend = j + 1; break; }
    }
    d_u += (1.1f * p.dmax) * T;
    float res = depth_obs - d_u;
    res = fminf(fmaxf(res, -0.30f), 0.30f);
    // pass 2 (backwards): acc[k] <- sum_{j>=k} acc_j
    float suffix = 0.f;
    for (int k = jend - 1; k >= 0; --k) {
        suffix += acc[k];
        acc[k] = suffix;
    }
    const float delta_d = (p.dmax - p.dmin) / (float)(D - 1);
    const float do_ds = -1.f / (2.f * p.th);
    int n = 0;
    for (int k = 0; k < jend; ++k) {
        float s = sdf[k * stride];
        if (s > -p.th && s < p.th) {
            float o = 0.5f - s / (2.f * p.th);
            float de_do = acc[k] / (1.f - o);
            if (de_do > 1e-2f) {
                emit(k, de_do * delta_d * do_ds, res);
                ++n;
            }
        }
    }
    return n;
}

template <class Emit>
DSP_HD int ray_scan(const float* sdf, int stride, float depth_obs, const RayParams& p, Emit emit) {
    float acc[DSP_MAX_D];
    return ray_scan_scratch(sdf, stride, depth_obs, p, acc, 1, emit);
}

}  // namespace dsp

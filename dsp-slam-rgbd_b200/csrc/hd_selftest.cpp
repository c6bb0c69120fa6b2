// Host build of hd_math.h for CPU-side unit tests (tests/test_hd_math.py).  TEST ONLY:
// not linked into libdspslam_sm100.so and never used by the product path.
#include "hd_math.h"

extern "C" {
void hd_exp_sim3(const float* x, float* T) { dsp::exp_sim3(x, T); }
void hd_exp_se3(const float* x, float* T) { dsp::exp_se3(x, T); }
void hd_inv_affine(const float* T, float* Ti) { dsp::inv_affine(T, Ti); }
void hd_mul_affine(const float* A, const float* B, float* C) { dsp::mul_affine(A, B, C); }
float hd_huber(float r, float b) { return dsp::huber_res(r, b); }
float hd_rot_prior(const float* T_co, float scale, float* J3) { return dsp::rot_prior(T_co, scale, J3); }
float hd_linspace_at(float a, float b, int n, int i) { return dsp::linspace_at(a, b, n, i); }
// returns the number of selected samples; idx/de_ds/res must hold D entries
int hd_ray_scan(const float* sdf, int D, float depth_obs, float th, float dmin, float dmax, int* idx,
                float* de_ds, float* res) {
    dsp::RayParams p;
    p.th = th; p.dmin = dmin; p.dmax = dmax; p.D = D; p.dtab = nullptr;
    int n = 0;
    return dsp::ray_scan(sdf, 1, depth_obs, p, [&](int k, float a, float r) {
        idx[n] = k; de_ds[n] = a; res[n] = r; ++n;
    });
}
}

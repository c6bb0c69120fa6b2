/* A plain C client of libdspslam_sm100.so: proves that include/dspslam.h is a C header (no C++ / torch types) and that
 * the boundary works without Python.  Built by tests/test_c_client.py with gcc; on a GPU box it creates a handle from
 * synthetic weights, evaluates the SDF and its Jacobian at a few points in the FP32 mode and in the tensor-core modes
 * and checks that they agree (1e-3) and that the Jacobian matches central differences of the SDF (FP32 mode).
 * usage: abi_client            (exit code 0 = ok, 77 = no CUDA device) */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "dspslam.h"

#define L 64
#define IN (L + 3)
#define N 1000

static float frand(unsigned* s) { *s = *s * 1664525u + 1013904223u; return ((*s >> 8) & 0xffff) / 65535.0f - 0.5f; }

int main(void) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { printf("no CUDA device\n"); return 77; }
    /* effective weights of a 9-layer decoder: [512,67] [512,512]x2 [445,512] [512,512]x4 [1,512] */
    const int out_dim[9] = {512, 512, 512, 512 - IN, 512, 512, 512, 512, 1};
    const int in_dim[9] = {IN, 512, 512, 512, 512, 512, 512, 512, 512};
    float* W[9]; float* b[9];
    unsigned seed = 7;
    for (int k = 0; k < 9; ++k) {
        W[k] = (float*)malloc(sizeof(float) * out_dim[k] * in_dim[k]);
        b[k] = (float*)malloc(sizeof(float) * out_dim[k]);
        const float sc = 2.4f / sqrtf((float)in_dim[k]);
        for (int i = 0; i < out_dim[k] * in_dim[k]; ++i) W[k][i] = sc * frand(&seed);
        for (int i = 0; i < out_dim[k]; ++i) b[k][i] = 0.05f * frand(&seed);
    }
    dspslam_decoder_desc desc = {L, 512, 8, 4};
    dspslam_handle* h = NULL;
    if (dspslam_create(&desc, (const float* const*)W, (const float* const*)b, 0, &h) != DSPSLAM_OK) {
        printf("create failed: %s\n", dspslam_last_error());
        return 1;
    }
    if (dspslam_abi_version() != DSPSLAM_ABI_VERSION) { printf("ABI mismatch\n"); return 1; }
    if (!dspslam_mode_available(h, DSPSLAM_MODE_TC)) { printf("tc unavailable: %s\n", dspslam_last_error()); return 1; }

    static float code[L], xyz[3 * (N + 6)], sdf[4][N + 6], jac[2][(N + 6) * IN];
    for (int i = 0; i < L; ++i) code[i] = 0.3f * frand(&seed);
    for (int i = 0; i < 3 * N; ++i) xyz[i] = 1.4f * frand(&seed);
    /* six extra points: +-eps around point 0 along x, y, z for the finite-difference check */
    const float eps = 1e-3f;
    for (int a = 0; a < 3; ++a)
        for (int sgn = 0; sgn < 2; ++sgn) {
            float* p = xyz + 3 * (N + 2 * a + sgn);
            memcpy(p, xyz, 3 * sizeof(float));
            p[a] += sgn ? eps : -eps;
        }
    float *d_code, *d_xyz, *d_sdf, *d_jac;
    cudaMalloc((void**)&d_code, sizeof code); cudaMalloc((void**)&d_xyz, sizeof xyz);
    cudaMalloc((void**)&d_sdf, sizeof sdf[0]); cudaMalloc((void**)&d_jac, sizeof jac[0]);
    cudaMemcpy(d_code, code, sizeof code, cudaMemcpyHostToDevice);
    cudaMemcpy(d_xyz, xyz, sizeof xyz, cudaMemcpyHostToDevice);
    const int modes[4] = {DSPSLAM_MODE_FP32, DSPSLAM_MODE_TC, DSPSLAM_MODE_TC_FAST, DSPSLAM_MODE_TC_EXACT};
    for (int m = 0; m < 4; ++m) {
        int rc = (m == 0 || m == 3) ? dspslam_sdf_jac(h, NULL, modes[m], d_code, d_xyz, N + 6, d_sdf, d_jac)
                                    : dspslam_sdf(h, NULL, modes[m], d_code, d_xyz, N + 6, d_sdf);
        if (rc != DSPSLAM_OK) { printf("mode %d failed: %s\n", modes[m], dspslam_last_error()); return 1; }
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error in mode %d\n", modes[m]); return 1; }
        cudaMemcpy(sdf[m], d_sdf, sizeof sdf[0], cudaMemcpyDeviceToHost);
        if (m == 0 || m == 3) cudaMemcpy(jac[m ? 1 : 0], d_jac, sizeof jac[0], cudaMemcpyDeviceToHost);
    }
    double e_tc = 0, e_fast = 0, e_exact = 0, e_jx = 0, e_fd = 0;
    for (int i = 0; i < N; ++i) {
        e_tc = fmax(e_tc, fabs(sdf[1][i] - sdf[0][i]));
        e_fast = fmax(e_fast, fabs(sdf[2][i] - sdf[0][i]));
        e_exact = fmax(e_exact, fabs(sdf[3][i] - sdf[0][i]));
    }
    int kinks = 0;
    for (int i = 0; i < N; ++i) {
        double row = 0;
        for (int c = 0; c < IN; ++c) row = fmax(row, fabs(jac[1][i * IN + c] - jac[0][i * IN + c]));
        if (row > 1e-4) ++kinks; else e_jx = fmax(e_jx, row);
    }
    for (int a = 0; a < 3; ++a) {
        const double fd = (sdf[0][N + 2 * a + 1] - sdf[0][N + 2 * a]) / (2.0 * eps);
        e_fd = fmax(e_fd, fabs(fd - jac[0][L + a]));
    }
    printf("sdf: tc %.2e  tc_fast %.2e  tc_exact %.2e | jacobian exact vs fp32 %.2e (%d kink rows) | d sdf/d xyz vs central differences %.2e\n",
           e_tc, e_fast, e_exact, e_jx, kinks, e_fd);
    const int ok = e_tc < 1e-3 && e_fast < 1e-3 && e_exact < 2e-5 && e_jx < 1e-4 && kinks <= 8 && e_fd < 2e-2;
    dspslam_destroy(h);
    printf(ok ? "c client ok\n" : "c client FAILED\n");
    return ok ? 0 : 1;
}

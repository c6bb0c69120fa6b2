// Shared device-side declarations of libdspslam_sm100 (B200 / sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "hd_math.h"

namespace dsp {

constexpr int HID = 512;             // hidden width
constexpr int MAX_L = 64;            // max code length
constexpr int POSE = 7;              // sim3 tangent dim
constexpr int HMAX = POSE + MAX_L;   // 71
constexpr int GROUP_ROWS = 128;      // rows per work unit (one partial-H slot, one tcgen05 M tile)
constexpr int PART_STRIDE = 5120;    // floats per partial slot: H (<=71x71) | b (<=71) | 2 loss sums
constexpr int PART_LOSS = 5112;      // [5112] sum rt^2 surface rows, [5113] sum rt^2 render rows
constexpr int REFINE_ROWS = 64;      // real rows per tile of the refinement kernel (hi rows | lo rows stacked to M = 128)
constexpr int MAX_PHASES = 4;        // front-to-back passes over the ray samples per iteration (k_gn_sample)
constexpr float SDF_OUTSIDE = 1e30f; // sentinel SDF of ray samples outside the unit ball (occupancy 0)

// Decoder parameters resident in HBM.  FP32 master copies in two layouts:
//  * "t" (k-major, [k_in][n_out]) for the forward contraction  out[n] = sum_k in[k] W[n][k]
//  * "n" (natural nn.Linear [out][in]) for the input-gradient sweep  g_in[i] = sum_o g[o] W[o][i]
// Layer 3 emits h3 = 512-(L+3) features, padded with zeros to h3p (multiple of 16).
struct DecoderDev {
    int L;       // code length
    int in_dim;  // L + 3
    int h3;      // 512 - in_dim      (445 for L = 64)
    int h3p;     // h3 rounded up to a multiple of 16 (448)
    // forward (transposed) ------------------------------------------------
    const float* Wt0;    // [in_dim][512]   rows 0..L-1: code part, rows L..L+2: xyz part
    const float* W0x4;   // [4][512]        xyz rows of layer 0 + one zero row (K padded to 4)
    const float* Wt[8];  // Wt[1],Wt[2]: [512][512]; Wt[3]: [512][h3p]; Wt[4]: [h3p][512] (hidden part,
                         // zero rows >= h3); Wt[5..7]: [512][512]; Wt[0] unused
    const float* W4in;   // [in_dim][512]   layer-4 weights of the re-injected input (code rows, then xyz rows)
    const float* bias[8];// bias[3] padded to h3p with zeros
    const float* w8;     // [512]
    float b8;
    // backward (natural) ---------------------------------------------------
    const float* Wn[8];  // Wn[1],Wn[2],Wn[5..7]: [512][512]; Wn[3]: [h3p][512] (zero rows >= h3);
                         // Wn[4]: [512][512], cols [0,h3) hidden | [h3,512) input; Wn[0] unused (Wt0 serves)
    // tcgen05 path (kernels_tc.cu); null until packed
    const uint8_t* tc_blob;  // FP16 weight stage images (forward program, then backward program)
    uint8_t* tc_xtra1;       // extras record (layers 0 and 4) of the single-code entry points
    float* tc_park;          // per-CTA scratch of the Jacobian kernels: layer-4 input gradients, [cta][128][68]
    uint4 tc_layers[2][8];   // per-layer stage table of the forward / backward sweep (kernels_tc.cu: LayerDesc)
};

// Per-object Gauss-Newton state (one per detection in the batch).
struct ObjState {
    float T_oc[16];   // camera -> object Sim(3) (current estimate)
    float T_co[16];   // its inverse (refreshed by k_prep every iteration)
    float scale, dmin, dmax, rot_res;
    float rotJ[3];
    float pad0;
    int status;       // DSPSLAM_ST_*
    int n_valid;      // ray samples inside the unit ball this iteration (evaluated or not)
    int K;            // render rows this iteration
    int M_active;     // surface rows in use (pose-only inlier cut)
    float sdf_loss, render_loss, loss;
    int rf_rows;      // ray samples of this iteration queued for the FP32-grade second evaluation (kernels_tc.cu: k_tc_jacx<2>)
    int ph_rows[MAX_PHASES];   // in-ball samples handed to the decoder in each ray phase of this iteration
};

struct JointParamsDev {
    float k1, k2, k3, k4, b1, b2, lr, s_damp, th;
    int D;
    int pose_only;    // 1: estimate_pose_cam_obj semantics (6 dof, raw residual, damping 1e-2)
};

// reference-shear / lattice voxel coordinates from the linear index (utils.py:97-116)
__device__ __forceinline__ void grid_coord(int64_t i, int G, int coord_mode, float& x, float& y, float& z) {
    const float vs = (float)(2.0 / (double)(G - 1));
    const float Gf = (float)G;
    const float zi = (float)(i % G);
    float xi, yi;
    if (coord_mode == 0) {
        const float q = __fdiv_rn(__ll2float_rn(i), Gf);
        yi = fmodf(q, Gf);
        xi = fmodf(__fdiv_rn(q, Gf), Gf);
    } else {
        yi = (float)((i / G) % G);
        xi = (float)((i / ((int64_t)G * G)) % G);
    }
    x = __fadd_rn(__fmul_rn(xi, vs), -1.f);
    y = __fadd_rn(__fmul_rn(yi, vs), -1.f);
    z = __fadd_rn(__fmul_rn(zi, vs), -1.f);
}

__device__ __forceinline__ int find_segment(const int* __restrict__ off, int n_seg, int idx) {
    // largest b with off[b] <= idx  (off has n_seg+1 entries, off[0] = 0)
    int lo = 0, hi = n_seg;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (off[mid] <= idx) lo = mid; else hi = mid;
    }
    return lo;
}

// x_obj = A * (ray * d) + t  in the reference's operation order (loss.py:74-77)
__device__ __forceinline__ void sample_point(const float* __restrict__ T, float rx, float ry, float rz,
                                             float d, float& x, float& y, float& z) {
    float px = rx * d, py = ry * d, pz = rz * d;
    x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(px, T[0]), __fmul_rn(py, T[1])), __fmul_rn(pz, T[2])), T[3]);
    y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(px, T[4]), __fmul_rn(py, T[5])), __fmul_rn(pz, T[6])), T[7]);
    z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(px, T[8]), __fmul_rn(py, T[9])), __fmul_rn(pz, T[10])), T[11]);
}

}  // namespace dsp

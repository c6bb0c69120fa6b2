// FP32 (FFMA) decoder forward and input-gradient sweep for a 32-row tile held in
// shared memory: the "FP32 validation mode" of the library (1e-5 parity with the
// reference's FP32 PyTorch path).  The tcgen05 path (mlp_tc.cuh) has the same
// structure with FP16 operands on the tensor cores.
//
//   forward  (deep_sdf/deep_sdf_decoder.py:75-110):  9 layers, ReLU, final tanh;
//            the latent code only enters as a per-object bias at layers 0 and 4
//            (c0 = b0 + W0[:, :L] z,  c4 = b4 + W4[:, h3:h3+L] z), so a tile only
//            carries xyz.
//   backward (what autograd does for loss_utils.py:82-113): g <- (g . relu') W_k,
//            k = 7..0; at layer 4 the 512-wide result splits into the hidden part
//            (continues) and the re-injected-input part (added at the end).
#pragma once
#include "common.cuh"

namespace dsp {
namespace simt {

constexpr int TM = 32;    // rows per tile
constexpr int NT = 256;   // threads per CTA: 4 row groups x 64 column groups, 8x8 outputs per thread
constexpr int LDA = 516;  // activation row stride in floats (512 + 4: keeps float4 alignment)
constexpr int LDJ = 72;   // row stride of the small per-row buffers

struct __align__(16) Smem {
    float act[2][TM * LDA];
    float xyz0[TM][4];       // x, y, z, 0
    float gin[TM][LDJ];      // input-part gradient from layer 4; later the J row (pose 7 | code L)
    float jac[TM][LDJ];      // d sdf / d [code | xyz]
    float sdf[TM];
    float dy[TM];            // 1 - tanh^2
    float coef[TM];          // de/ds of the row (1 for surface rows)
    float res[TM];           // precomputed residual (render rows)
    float wrow[TM];          // normal-equation weight of the row
    float rt[TM];            // robust residual
    uint8_t mask[8][TM][64]; // ReLU sign bits, 8 columns per byte
};

// out[32 x N] = in[32 x K] * Wt[K x N]; each thread owns rows rg*8.. and columns cg*8..;
// `epi(row, n0, acc[8])` consumes 8 consecutive columns of one row.
template <class Epi>
__device__ __forceinline__ void gemm32(const float* in, int lda, int K, const float* __restrict__ Wt,
                                       int ldw, int N, Epi epi) {
    const int tid = threadIdx.x;
    const int cg = tid & 63, rg = tid >> 6;
    const int n0 = cg * 8;
    if (n0 >= N) return;
    float acc[8][8];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = 0.f;
    const float* wp = Wt + n0;
    const float* ap = in + rg * 8 * lda;
    float4 w[4][2];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        const float4* p = reinterpret_cast<const float4*>(wp + (size_t)kk * ldw);
        w[kk][0] = __ldg(p);
        w[kk][1] = __ldg(p + 1);
    }
    for (int k = 0; k < K; k += 4) {
        float4 wn[4][2];
        const bool more = (k + 4 < K);
        if (more) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const float4* p = reinterpret_cast<const float4*>(wp + (size_t)(k + 4 + kk) * ldw);
                wn[kk][0] = __ldg(p);
                wn[kk][1] = __ldg(p + 1);
            }
        }
        float4 a[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) a[r] = *reinterpret_cast<const float4*>(ap + r * lda + k);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const float av = (kk == 0) ? a[r].x : (kk == 1) ? a[r].y : (kk == 2) ? a[r].z : a[r].w;
                acc[r][0] = fmaf(av, w[kk][0].x, acc[r][0]);
                acc[r][1] = fmaf(av, w[kk][0].y, acc[r][1]);
                acc[r][2] = fmaf(av, w[kk][0].z, acc[r][2]);
                acc[r][3] = fmaf(av, w[kk][0].w, acc[r][3]);
                acc[r][4] = fmaf(av, w[kk][1].x, acc[r][4]);
                acc[r][5] = fmaf(av, w[kk][1].y, acc[r][5]);
                acc[r][6] = fmaf(av, w[kk][1].z, acc[r][6]);
                acc[r][7] = fmaf(av, w[kk][1].w, acc[r][7]);
            }
        }
        if (more) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                w[kk][0] = wn[kk][0];
                w[kk][1] = wn[kk][1];
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) epi(rg * 8 + r, n0, acc[r]);
}

__device__ __forceinline__ void store8(float* dst, const float* v) {
    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// Forward pass of the tile whose coordinates sit in sm.xyz0.  Leaves sdf in sm.sdf,
// 1-sdf^2 in sm.dy and (if JAC) the ReLU masks in sm.mask.  Ends with __syncthreads().
template <bool JAC>
__device__ void forward32(Smem& sm, const DecoderDev& dec, const float* __restrict__ c0,
                          const float* __restrict__ c4) {
    float* A = sm.act[0];
    float* B = sm.act[1];
    // bias + ReLU epilogue shared by all hidden layers
    auto relu_epi = [&](float* out, const float* __restrict__ bias, int layer) {
        return [&sm, out, bias, layer](int row, int n0, float* acc) {
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + n0));
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + n0) + 1);
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            float v[8];
            uint32_t m = 0;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float t = acc[c] + bb[c];
                m |= (t > 0.f ? 1u : 0u) << c;
                v[c] = fmaxf(t, 0.f);
            }
            store8(out + row * LDA + n0, v);
            if (JAC) sm.mask[layer][row][n0 >> 3] = (uint8_t)m;
        };
    };
    // layer 0: K = 3 (+1 zero) from xyz, latent part folded into c0
    gemm32(&sm.xyz0[0][0], 4, 4, dec.W0x4, HID, HID, relu_epi(A, c0, 0));
    __syncthreads();
    gemm32(A, LDA, HID, dec.Wt[1], HID, HID, relu_epi(B, dec.bias[1], 1));
    __syncthreads();
    gemm32(B, LDA, HID, dec.Wt[2], HID, HID, relu_epi(A, dec.bias[2], 2));
    __syncthreads();
    gemm32(A, LDA, HID, dec.Wt[3], dec.h3p, dec.h3p, relu_epi(B, dec.bias[3], 3));
    __syncthreads();
    {   // layer 4: hidden part (K = h3p) + xyz part + latent bias c4
        const float* __restrict__ W4x = dec.W4in + (size_t)dec.L * HID;
        gemm32(B, LDA, dec.h3p, dec.Wt[4], HID, HID, [&sm, A, c4, W4x](int row, int n0, float* acc) {
            const float x = sm.xyz0[row][0], y = sm.xyz0[row][1], z = sm.xyz0[row][2];
            float v[8];
            uint32_t m = 0;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int n = n0 + c;
                float t = acc[c];
                t = fmaf(x, __ldg(W4x + n), t);
                t = fmaf(y, __ldg(W4x + HID + n), t);
                t = fmaf(z, __ldg(W4x + 2 * HID + n), t);
                t += __ldg(c4 + n);
                m |= (t > 0.f ? 1u : 0u) << c;
                v[c] = fmaxf(t, 0.f);
            }
            store8(A + row * LDA + n0, v);
            if (JAC) sm.mask[4][row][n0 >> 3] = (uint8_t)m;
        });
    }
    __syncthreads();
    gemm32(A, LDA, HID, dec.Wt[5], HID, HID, relu_epi(B, dec.bias[5], 5));
    __syncthreads();
    gemm32(B, LDA, HID, dec.Wt[6], HID, HID, relu_epi(A, dec.bias[6], 6));
    __syncthreads();
    gemm32(A, LDA, HID, dec.Wt[7], HID, HID, relu_epi(B, dec.bias[7], 7));
    __syncthreads();
    // layer 8 (1 x 512) + tanh: warp w handles rows 4w .. 4w+3
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
        const int row = warp * 4 + rr;
        float s = 0.f;
        for (int n = lane; n < HID; n += 32) s = fmaf(B[row * LDA + n], __ldg(dec.w8 + n), s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) {
            const float y = tanhf(s + dec.b8);
            sm.sdf[row] = y;
            sm.dy[row] = 1.f - y * y;
        }
    }
    __syncthreads();
}

// Input-gradient sweep; needs sm.mask / sm.dy from forward32<true>.  Leaves
// d sdf / d [code | xyz] in sm.jac[row][0 .. in_dim).  Ends with __syncthreads().
__device__ void backward32(Smem& sm, const DecoderDev& dec) {
    float* A = sm.act[0];
    float* B = sm.act[1];
    const int tid = threadIdx.x;
    // g7 = dy * w8 . relu'(a7)  -> A
    for (int e = tid; e < TM * 64; e += NT) {
        const int row = e >> 6, byte = e & 63;
        const uint32_t m = sm.mask[7][row][byte];
        const float dy = sm.dy[row];
        float v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) v[c] = ((m >> c) & 1u) ? dy * __ldg(dec.w8 + byte * 8 + c) : 0.f;
        store8(A + row * LDA + byte * 8, v);
    }
    __syncthreads();
    auto mask_epi = [&](float* out, int layer_below) {
        return [&sm, out, layer_below](int row, int n0, float* acc) {
            const uint32_t m = sm.mask[layer_below][row][n0 >> 3];
            float v[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = ((m >> c) & 1u) ? acc[c] : 0.f;
            store8(out + row * LDA + n0, v);
        };
    };
    gemm32(A, LDA, HID, dec.Wn[7], HID, HID, mask_epi(B, 6));
    __syncthreads();
    gemm32(B, LDA, HID, dec.Wn[6], HID, HID, mask_epi(A, 5));
    __syncthreads();
    gemm32(A, LDA, HID, dec.Wn[5], HID, HID, mask_epi(B, 4));
    __syncthreads();
    {   // layer 4: columns [0,h3) continue (masked by layer 3), [h3,512) are d/d(input)
        const int h3 = dec.h3, h3p = dec.h3p;
        gemm32(B, LDA, HID, dec.Wn[4], HID, HID, [&sm, A, h3, h3p](int row, int n0, float* acc) {
            const uint32_t m = sm.mask[3][row][n0 >> 3];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int n = n0 + c;
                if (n < h3) {
                    A[row * LDA + n] = ((m >> c) & 1u) ? acc[c] : 0.f;
                } else {
                    if (n < h3p) A[row * LDA + n] = 0.f;
                    sm.gin[row][n - h3] = acc[c];
                }
            }
        });
    }
    __syncthreads();
    gemm32(A, LDA, dec.h3p, dec.Wn[3], HID, HID, mask_epi(B, 2));
    __syncthreads();
    gemm32(B, LDA, HID, dec.Wn[2], HID, HID, mask_epi(A, 1));
    __syncthreads();
    gemm32(A, LDA, HID, dec.Wn[1], HID, HID, mask_epi(B, 0));
    __syncthreads();
    // layer 0: jac[row][i] = gin[row][i] + sum_o g0[row][o] W0[o][i]   (Wt0 is [i][o])
    const int in_dim = dec.in_dim;
    for (int e = tid; e < TM * in_dim; e += NT) {
        const int row = e / in_dim, i = e - row * in_dim;
        const float4* g = reinterpret_cast<const float4*>(B + row * LDA);
        const float4* w = reinterpret_cast<const float4*>(dec.Wt0 + (size_t)i * HID);
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 4
        for (int o = 0; o < HID / 4; ++o) {
            const float4 gv = g[o];
            const float4 wv = __ldg(w + o);
            s0 = fmaf(gv.x, wv.x, s0);
            s1 = fmaf(gv.y, wv.y, s1);
            s2 = fmaf(gv.z, wv.z, s2);
            s3 = fmaf(gv.w, wv.w, s3);
        }
        sm.jac[row][i] = (s0 + s1) + (s2 + s3) + sm.gin[row][i];
    }
    __syncthreads();
}

}  // namespace simt
}  // namespace dsp

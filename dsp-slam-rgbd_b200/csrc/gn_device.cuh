// Device-side pieces shared by the Gauss-Newton glue kernels (kernels_simt.cu) and the tcgen05 decoder
// kernels (kernels_tc.cu): the per-object "extra block" images of layers 0 and 4 and the 128-row tile
// work lists.  Both used to be kernels of their own; they are now executed by the kernel that produces
// their inputs (k_gn_prep / k_gn_solve, k_gn_sample / k_gn_rays), which removes five launches per iteration.
#pragma once
#include "kernels.h"

namespace dsp {

constexpr int XTRA_BYTES = 16384;  // one extra-block B image: 512 rows x 32 B

__device__ __forceinline__ void split_f16(float v, __half& hi, __half& lo) {
    hi = __float2half_rn(v);
    lo = __float2half_rn(v - __half2float(hi));
}

// Extra block of layer 0 (which = 0) or 4 (which = 1), output n: the B operand of the K=16 MMA that applies the
// bias, the per-object latent term and the xyz columns (kernels_tc.cu):
//   k=0,1: hi/lo of the latent bias c;  k=2..4: hi(W_xyz) (x x_hi);  k=5..7: hi(W_xyz) (x x_lo);  k=8..10: lo(W_xyz) (x x_hi)
// image layout (no swizzle, K-major): byte(n, k) = (n/8)*256 + (k/8)*128 + (n%8)*16 + (k%8)*2
__device__ __forceinline__ void pack_extra_row(const DecoderDev& dec, int which, int n, float bias, uint8_t* image) {
    float wx[3];
#pragma unroll
    for (int c = 0; c < 3; ++c)
        wx[c] = which ? __ldg(dec.W4in + (size_t)(dec.L + c) * HID + n) : __ldg(dec.W0x4 + (size_t)c * HID + n);
    __align__(16) __half h[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) h[k] = __float2half_rn(0.f);
    split_f16(bias, h[0], h[1]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        __half hi, lo;
        split_f16(wx[c], hi, lo);
        h[2 + c] = hi; h[5 + c] = hi; h[8 + c] = lo;
    }
    uint8_t* dst = image + (size_t)(n >> 3) * 256 + (size_t)(n & 7) * 16;
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(&h[0]);
    *reinterpret_cast<uint4*>(dst + 128) = *reinterpret_cast<const uint4*>(&h[8]);
}

// rows of the object's Jacobian list that lie in tiles made of surface rows only: known as soon as the iteration's
// state is, so their tiles run on a second stream beside the ray pipeline (GnWork::early)
__device__ __forceinline__ int early_rows(const GnWork& w, int b) {
    return w.early ? ((w.pts_off[b + 1] - w.pts_off[b]) / w.jac_rows) * w.jac_rows : 0;
}

// rows object b contributes to work list `which` (0: ray samples of the current phase, 1: Jacobian rows (behind the
// early tiles), 2: ray samples queued for the second, FP32-grade evaluation - tiles of REFINE_ROWS rows, 3: the
// early Jacobian tiles)
// (volatile: the fields were written by other CTAs of the running kernel; an L1 line cached earlier by this CTA
// could be stale)
__device__ __forceinline__ int plan_rows(const GnWork& w, int which, int b) {
    const volatile ObjState* st = w.st + b;
    if (st->status != 0) return 0;
    if (which == 0) return st->ph_rows[w.phase];
    if (which == 2) return st->rf_rows;
    if (which == 3) return early_rows(w, b);
    return (w.pts_off[b + 1] - w.pts_off[b]) + (w.prm.pose_only ? 0 : st->K) - early_rows(w, b);
}

// Work list of the tcgen05 kernels: one entry (object, first row) per 128-row tile, objects in order.
// Executed by ONE whole CTA (any block size that is a multiple of 32, <= 1024).
//   pass 1: tile_off[b] = exclusive scan of the per-object tile counts (global scratch, B+1 ints)
//   pass 2: every thread fills tiles t = tid, tid + blockDim, ...: object by binary search in tile_off
static __device__ void plan_tiles(const GnWork& w, int which) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    __shared__ int s_off[1025];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
    const int TR = (which == 2) ? REFINE_ROWS : (which == 0 ? GROUP_ROWS : w.jac_rows);      // rows per tile
    int* out_obj = (which == 3) ? w.etile_obj : w.tile_obj;
    int* out_row0 = (which == 3) ? w.etile_row0 : w.tile_row0;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < w.B; base += blockDim.x) {
        const int b = base + tid;
        const int nt = (b < w.B) ? (plan_rows(w, which, b) + TR - 1) / TR : 0;
        int inc = nt;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            const int ws = (lane < nw) ? s_warp[lane] : 0;
            int winc = ws;
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += t;
            }
            s_warp[lane] = winc - ws;                   // exclusive warp offsets
        }
        __syncthreads();
        const int first = s_carry + s_warp[wid] + inc - nt;
        if (b < w.B) {
            w.tile_off[b] = first;
            if (b < 1024) s_off[b] = first;
        }
        __syncthreads();
        if (tid == blockDim.x - 1) s_carry = first + nt;
        __syncthreads();
    }
    const int total = s_carry;
    if (tid == 0) {
        w.tile_off[w.B] = total;
        if (w.B <= 1024) s_off[w.B] = total;
        w.n_tiles[which] = total;
    }
    __threadfence_block();
    __syncthreads();
    const int* off = (w.B <= 1024) ? s_off : w.tile_off;   // (global: written above by this CTA, fenced)
    for (int t = tid; t < total; t += blockDim.x) {
        const int b = find_segment(off, w.B, t);
        out_obj[t] = b;
        out_row0[t] = (t - off[b]) * TR + (which == 1 ? early_rows(w, b) : 0);
    }
}

// "last CTA done" election: every CTA of the grid calls this once after its global writes; returns true
// in exactly one CTA (all its threads), after which that CTA sees the writes of all the others.
__device__ __forceinline__ bool last_cta_done(unsigned int* counter) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned total = gridDim.x * gridDim.y;
        const unsigned ticket = atomicAdd(counter, 1u);
        s_last = (ticket == total - 1);
        if (s_last) *counter = 0;                        // re-arm for the next launch on this stream
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0;
}

}  // namespace dsp

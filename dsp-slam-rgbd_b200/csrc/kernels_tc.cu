// tcgen05 decoder kernels of libdspslam_sm100 (DSPSLAM_MODE_TC).
//
// One persistent, warp-specialised kernel template evaluates the DeepSDF decoder
// (deep_sdf/deep_sdf_decoder.py:75-110) - and, for the Jacobian variants, the input-gradient
// sweep autograd performs for loss_utils.py:82-113 - on 128-row tiles:
//
//   * every 512-wide layer is a point x weight contraction on the 5th-generation tensor cores:
//     D[128 x N] (FP32, TMEM) += A[128 x K] (FP16, shared memory) * W^T[K x N] (FP16, shared memory),
//     issued by one thread as M128 x N<=256 x K16 `tcgen05.mma.cta_group::1.kind::f16`;
//   * the weights are pre-packed on the host into 16 KB stage images (256 outputs x 32 K values) that are
//     already in the 64-byte-swizzled K-major layout the MMA reads, and are streamed from L2 by a producer
//     thread with 1-D bulk async copies (TMA) through a ring of mbarrier-guarded stages;
//   * activations never leave the SM: the epilogue warps read the accumulators from TMEM
//     (`tcgen05.ld`), apply ReLU while converting to FP16 and write the next layer's A operand
//     straight into the swizzled shared-memory tile; ReLU sign bits are kept per thread for the
//     backward sweep, which runs the transposed weight images through the same pipeline;
//   * biases, the per-object latent contribution (c0/c4, see k_latent_bias) and the xyz columns
//     of layers 0 and 4 enter as one extra K=16 MMA per layer: A_x = [1 1 x_hi x_lo ...],
//     B_x = [b_hi b_lo W_hi W_hi W_lo ...] (hi/lo FP16 splits, so these terms keep ~FP32 accuracy);
//   * the GN variants read their rows from the work lists built by the mode-independent kernels
//     (kernels_simt.cu) and, for the Jacobian variant, reduce J^T J / J^T r per tile in shared memory.
#include "kernels_tc.h"

#include <cstring>
#include <string>
#include <vector>

#include "gn_device.cuh"
#include "tc_ptx.cuh"

namespace dsp {
namespace tc {

using namespace ptx;

constexpr int TILE_M = 128;
constexpr int NT = 320;            // warps 0-7: epilogue, warp 8: weight producer, warp 9: MMA issuer
constexpr int EPI_THREADS = 256;
constexpr int A_BLK = 16384;       // one K=64 block of the activation tile: 128 rows x 128 B, SW128
constexpr int A_BYTES = 8 * A_BLK;
constexpr int AX_BYTES = 4096;     // the K=16 extra block of the tile: 128 rows x 32 B, no swizzle
constexpr int STAGE_BYTES = 16384;
constexpr int JLD = 76;            // row stride (floats) of the per-row Jacobian buffer (float4-aligned rows)
constexpr int PARK_LD = 68;        // columns of the layer-4 input-gradient park in global memory, per CTA [PARK_LD][TILE_M]:
                                   // column-major, so that the 32 rows of a warp store / load one 128-byte line
constexpr int PARK_CTAS = 256;     // CTAs the park is sized for (grids are capped at the SM count)
constexpr float GSCALE = 256.f;    // gradients are carried scaled so that FP16 keeps them normal
constexpr int TMEM_COLS = 512;

// What the producer and the MMA warp walk: one entry per layer of the forward ([0]) and backward ([1])
// sweep.  The hidden stages of a layer (n in chunks of 256 outputs x k32 chunks of 32 K values, each a
// 64B-swizzled image of min(256, n - n0) * 64 bytes) lie contiguously in the blob in consumption order.
struct __align__(16) LayerDesc {
    uint32_t src;    // blob offset of the first hidden stage
    uint32_t xsrc;   // offset of the K=16 extra block (xkind 1: in the blob, 2: in the object's extras record)
    uint16_t n;      // accumulator columns
    uint8_t k32;     // 32-wide K chunks of the hidden part (0: extras only)
    uint8_t xkind;   // 0: no extra block
    uint32_t src_lo; // forward sweep: blob offset of the same stage sequence built from lo(W) = fp16(W - fp16(W))
};
static_assert(sizeof(LayerDesc) == 16, "LayerDesc layout");
static_assert(sizeof(LayerDesc) == sizeof(uint4), "DecoderDev::tc_layers holds LayerDesc entries");
// The table travels with the decoder view (DecoderDev::tc_layers, kernel parameter space): it belongs to the handle,
// so two handles holding different decoders on one device cannot overwrite each other's table.

struct TcArgs {
    DecoderDev dec;
    const uint8_t* blob;     // weight stage images
    const uint8_t* xtra;     // per-object extras: [obj][2][XTRA_BYTES] (layer 0, layer 4)
    // stand-alone row sources / sinks
    const float* xyz;
    int64_t n, i0;
    int G;
    float* out;
    float* jac_out;
    int n_tiles;
    int park_cta0;           // first per-CTA park block of this launch (two Jacobian launches may run side by side)
    int wsplit;              // bit l set: the forward sweep of the Jacobian variants issues every hidden K chunk of layer l
                             // against hi(W) and lo(W) (DSPSLAM_MODE_TC)
    // Gauss-Newton work lists
    GnWork w;
};

#ifndef DSP_TC_PAIR
#define DSP_TC_PAIR 1
#endif

template <int MODE>
struct Cfg {
    static constexpr bool JAC = MODE >= 4;
    static constexpr bool GN = (MODE == 3 || MODE == 5);
    static constexpr int NST = 5;
    // the kernel runs as clusters of two CTAs that walk the same weight sequence: every hidden stage is fetched from
    // L2 once per pair (each CTA issues one half, multicast to both), halving the L2 -> SM traffic
    // (DSP_TC_PAIR: 0 = never, 1 = forward-only variants, 2 = all variants)
    static constexpr bool PAIR = DSP_TC_PAIR == 2 || (DSP_TC_PAIR == 1 && !JAC);
    static constexpr int O_AX = A_BYTES;
    static constexpr int O_ST = O_AX + AX_BYTES;
    static constexpr int O_W8 = O_ST + NST * STAGE_BYTES;
    static constexpr int O_XYZ = O_W8 + 2048;
    static constexpr int O_PART = O_XYZ + 2048;
    static constexpr int O_DY = O_PART + 1024;
    static constexpr int O_BAR = O_DY + 512;
    // the per-row Jacobian buffer (TILE_M x JLD floats) lives at offset 0, over the activation tile, which is dead
    // once the last backward MMA has completed and is not written again before the tile's J^T J is done
    static constexpr int O_WROW = O_BAR + 256;
    static constexpr int O_RT = O_WROW + (JAC ? 512 : 0);
    static constexpr int O_END = O_RT + (JAC ? 512 : 0);
    static constexpr int SMEM = O_END + 1024;   // + alignment slack
};

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__device__ __forceinline__ void build_pose_row(float* J, float gx, float gy, float gz, float x, float y, float z) {
    // d x_o / d xi = [I | -hat(x) | x]  (loss_utils.py:176-195)
    J[0] = gx; J[1] = gy; J[2] = gz;
    J[3] = y * gz - z * gy;
    J[4] = z * gx - x * gz;
    J[5] = x * gy - y * gx;
    J[6] = gx * x + gy * y + gz * z;
}

// MODE 0: rows = xyz array -> out            1: voxel grid, reference shear -> out   2: voxel grid, lattice -> out
//      3: GN forward (compacted ray samples) 4: xyz array -> sdf + jac               5: GN Jacobian rows -> J^T J partials
// -DDSP_TC_TRACE: CTA 0 of the GN Jacobian launch (MODE 5) logs (clock64, tag) pairs of its first epilogue warp and of
// its MMA warp into the unused tail of the park buffer (tools/trace_jac.py reads them): where a layer step's time goes.
#ifdef DSP_TC_TRACE
#define TRACE_E(tag) do { if (MODE == 5 && blockIdx.x == 0 && warp == 0 && lane == 0 && te < 3900) { trace[te++] = clock64(); trace[te++] = (tag); } } while (0)
#define TRACE_M(tag) do { if (MODE == 5 && blockIdx.x == 0 && lane == 0 && tm < 3900) { trace[4096 + tm++] = clock64(); trace[4096 + tm++] = (tag); } } while (0)
#else
#define TRACE_E(tag) do { } while (0)
#define TRACE_M(tag) do { } while (0)
#endif

template <int MODE>
__global__ void __launch_bounds__(NT, 1) k_tc(const __grid_constant__ TcArgs a) {
    using C = Cfg<MODE>;
#ifdef DSP_TC_TRACE
    long long* trace = reinterpret_cast<long long*>(a.dec.tc_park + (size_t)200 * TILE_M * PARK_LD);
    int te = 0, tm = 0;
#endif
    constexpr bool JAC = C::JAC;
    constexpr int NST = C::NST;
    constexpr bool PAIR = C::PAIR;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (sbase - raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    const uint32_t sA = sbase, sAX = sbase + C::O_AX, sST = sbase + C::O_ST;
    float* w8s = reinterpret_cast<float*>(sm + C::O_W8);
    float4* xyzs = reinterpret_cast<float4*>(sm + C::O_XYZ);
    float* parts = reinterpret_cast<float*>(sm + C::O_PART);
    float* dys = reinterpret_cast<float*>(sm + C::O_DY);
    const uint32_t bar0 = sbase + C::O_BAR;
    auto bar_full = [&](int s) { return bar0 + 8u * s; };
    auto bar_empty = [&](int s) { return bar0 + 8u * (NST + s); };
    const uint32_t bar_aready = bar0 + 8u * (2 * NST), bar_dfull = bar0 + 8u * (2 * NST + 1);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + C::O_BAR + 8 * (2 * NST + 2));
    float* jrow = reinterpret_cast<float*>(sm);
    static_assert(TILE_M * JLD * 4 <= A_BYTES, "Jacobian rows must fit in the activation tile");
    float* wrows = reinterpret_cast<float*>(sm + C::O_WROW);
    float* rts = reinterpret_cast<float*>(sm + C::O_RT);

    if (tid == 0) {
        for (int s = 0; s < NST; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), PAIR ? 2 : 1); }
        mbar_init(bar_aready, 8);
        mbar_init(bar_dfull, 1);
        fence_mbar_init();
    }
    if (warp == 9) {
        tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
        tmem_relinquish();
    }
    for (int i = tid; i < HID; i += NT) w8s[i] = __ldg(a.dec.w8 + i);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (PAIR) cluster_sync_all();                           // the peer's barriers exist before anything is sent to them
    const uint32_t tmem = *tmem_slot;

    int n_tiles = a.n_tiles;
    if (C::GN) n_tiles = a.w.n_tiles[MODE == 3 ? 0 : 1];
    constexpr int NPROG = JAC ? 2 : 1;
    // hi(W) + lo(W) passes in the FORWARD sweep only: the forward weights set the ReLU pattern and the SDF value the
    // Jacobian hangs on; rounding the backward sweep's weights moves a GN step by nothing measurable
    // (oracle/precision_study2.py: "J Wfwd a g" 98.8 % of steps = "J W a g"; "J Wbwd a g" 97.5 % = single pass)
    const int wsplit = JAC ? a.wsplit : 0;
    // tile walk: a pair takes tiles (tb, tb + 1); both CTAs make the same number of trips (the multicast protocol
    // needs them in step), a CTA whose tile index falls behind the list runs the trip with no valid rows
    const int rank = PAIR ? (int)(blockIdx.x & 1u) : 0;
    const int tb0 = (int)blockIdx.x - rank;

    if (warp == 8) {
        // ===================== weight producer (whole warp, one elected lane issues) =====================
        uint32_t s = 0, ph = 0;
        for (int tb = tb0; tb < n_tiles; tb += gridDim.x) {
            const int t = min(tb + rank, n_tiles - 1);
            const int obj = C::GN ? a.w.tile_obj[t] : 0;
            const uint8_t* xo = a.xtra + (size_t)obj * (2 * XTRA_BYTES);
            for (int p = 0; p < NPROG; ++p) {
#pragma unroll 1
                for (int li = 0; li < 8; ++li) {
                    const int nrep = (p == 0 && ((wsplit >> li) & 1)) ? 2 : 1;
                    const LayerDesc ld = *reinterpret_cast<const LayerDesc*>(&a.dec.tc_layers[p][li]);
                    if (ld.xkind) {                              // the K=16 extra block (bias / latent / xyz)
                        mbar_wait(bar_empty(s), ph ^ 1u);
                        if (elect_one()) {
                            const uint32_t bytes = (uint32_t)ld.n * 32u;
                            mbar_expect_tx(bar_full(s), bytes);
                            bulk_g2s(sST + s * STAGE_BYTES, (ld.xkind == 2 ? xo : a.blob) + ld.xsrc, bytes, bar_full(s));
                        }
                        __syncwarp();
                        if (++s == NST) { s = 0; ph ^= 1u; }
                    }
                    const uint8_t* src = a.blob + ld.src;        // hidden stages are contiguous in consumption order
                    const uint8_t* srcl = a.blob + ld.src_lo;    // the same sequence packed from lo(W)
                    for (int n0 = 0; n0 < ld.n; n0 += 256) {
                        const uint32_t bytes = (uint32_t)min(256, (int)ld.n - n0) * 64u;
#pragma unroll 1
                        for (int k2 = 0; k2 < nrep * ld.k32; ++k2) {
                            const uint8_t* from = (JAC && nrep == 2 && (k2 & 1)) ? srcl : src;
                            mbar_wait(bar_empty(s), ph ^ 1u);
                            if (elect_one()) {
                                mbar_expect_tx(bar_full(s), bytes);
                                if (PAIR) {
                                    const uint32_t half = bytes >> 1;
                                    bulk_g2s_mc(sST + s * STAGE_BYTES + rank * half, from + rank * half, half, bar_full(s), 3);
                                } else {
                                    bulk_g2s(sST + s * STAGE_BYTES, from, bytes, bar_full(s));
                                }
                            }
                            __syncwarp();
                            if (nrep == 1 || (k2 & 1)) { src += bytes; srcl += bytes; }
                            if (++s == NST) { s = 0; ph ^= 1u; }
                        }
                    }
                }
            }
        }
        if (PAIR) {                                         // tail: every arrive the peer still owes this CTA has landed
            for (int i = 0; i < NST; ++i) {
                mbar_wait(bar_empty(s), ph ^ 1u);
                if (++s == NST) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 9) {
        // ===================== MMA issuer (whole warp, one elected lane issues) =====================
        uint32_t s = 0, ph = 0, aph = 0;
        const uint64_t adesc0 = smem_desc(sA, 16, 1024, 2);      // activations: 128B swizzle, 8-row groups 1 KB apart
        const uint64_t bdesc0 = smem_desc(sST, 16, 512, 4);      // weight stage: 64B swizzle, 8-row groups 512 B apart
        const uint64_t axdesc = smem_desc(sAX, 128, 256, 0);
        const uint64_t bxdesc0 = smem_desc(sST, 128, 256, 0);
        auto release = [&](uint32_t bar) {                   // stage reusable (in both CTAs of a pair) once these MMAs finish
            if (PAIR) umma_commit_mc(bar, 3); else umma_commit(bar);
        };
        for (int tb = tb0; tb < n_tiles; tb += gridDim.x) {
            for (int p = 0; p < NPROG; ++p) {
#pragma unroll 1
                for (int li = 0; li < 8; ++li) {
                    const int nrep = (p == 0 && ((wsplit >> li) & 1)) ? 2 : 1;
                    const LayerDesc ld = *reinterpret_cast<const LayerDesc*>(&a.dec.tc_layers[p][li]);
                    mbar_wait(bar_aready, aph);                  // this layer's A operand written, accumulator drained
                    aph ^= 1u;
                    TRACE_M(100 * p + li);
                    if (ld.xkind) {
                        mbar_wait(bar_full(s), ph);
                        tc_fence_after();
                        if (elect_one()) {
                            const uint64_t bx = bxdesc0 + (uint64_t)((s * STAGE_BYTES) >> 4);
                            for (int n0 = 0; n0 < ld.n; n0 += 256) {
                                const int nn = min(256, (int)ld.n - n0);
                                umma_f16(tmem + n0, axdesc, bx + (uint64_t)((n0 * 32) >> 4), idesc_f16(TILE_M, nn), 0u);
                            }
                            release(bar_empty(s));
                        }
                        __syncwarp();
                        if (++s == NST) { s = 0; ph ^= 1u; }
                    }
                    for (int n0 = 0; n0 < ld.n; n0 += 256) {
                        const uint32_t id = idesc_f16(TILE_M, min(256, (int)ld.n - n0));
                        const uint32_t dcol = tmem + n0;
#pragma unroll 1
                        for (int k2 = 0; k2 < nrep * ld.k32; ++k2) {
                            const int kc = (nrep == 2) ? (k2 >> 1) : k2;     // (hi(W) stage, then lo(W) stage of the same chunk)
                            mbar_wait(bar_full(s), ph);
                            tc_fence_after();
                            if (elect_one()) {
                                // A: K chunk kc = block kc/2 (16 KB apart), half kc%2 (64 B apart); +32 B per K=16 step
                                const uint64_t ad = adesc0 + (uint64_t)((kc >> 1) * (A_BLK >> 4) + (kc & 1) * 4);
                                const uint64_t bd = bdesc0 + (uint64_t)((s * STAGE_BYTES) >> 4);
                                umma_f16(dcol, ad, bd, id, (ld.xkind == 0 && k2 == 0) ? 0u : 1u);
                                umma_f16(dcol, ad + 2, bd + 2, id, 1u);
                                release(bar_empty(s));
                            }
                            __syncwarp();
                            if (++s == NST) { s = 0; ph ^= 1u; }
                        }
                    }
                    if (elect_one()) umma_commit(bar_dfull);     // accumulator of the layer complete
                    __syncwarp();
                    TRACE_M(1000 + 100 * p + li);
                }
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int q = warp & 3, hc = warp >> 2;
        const int r = q * 32 + lane;                       // tile row = TMEM lane
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
        const int rx = r & 7;
        const uint32_t a_row = sA + (uint32_t)(r >> 3) * 1024u + (uint32_t)rx * 128u;
        const int etid = tid;                              // 0..255
        uint32_t dph = 0;
        uint32_t mw[8][8];                                 // ReLU sign words of this thread's 256 columns, per layer
        const int L = a.dec.L;
        // layer-4 input gradients of row r wait here (written and read back by the hc == 1 thread of the row)
        float* park = JAC ? a.dec.tc_park + (size_t)(blockIdx.x + a.park_cta0) * TILE_M * PARK_LD + r : nullptr;
#define PARK(c) park[(c) * TILE_M]

        auto signal_ready = [&]() {
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_aready);
        };
        // write 32 consecutive FP16 values (16 packed words) of row r at column `col` of the activation tile
        auto store32 = [&](int col, const uint32_t* p) {
            const uint32_t blk = a_row + (uint32_t)(col >> 6) * A_BLK;
            const int cc0 = (col & 63) >> 3;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq)
                st_shared_v4(blk + (uint32_t)(((cc0 + qq) ^ rx) << 4), p[4 * qq], p[4 * qq + 1], p[4 * qq + 2], p[4 * qq + 3]);
        };

        for (int tb = tb0; tb < n_tiles; tb += gridDim.x) {
            const bool idle = tb + rank >= n_tiles;        // (pairs only) the odd CTA's trip behind the last tile
            const int t = idle ? n_tiles - 1 : tb + rank;
            // ---------------- tile rows ----------------
            int obj = 0, row0 = 0, rows_valid = 0;
            int dest = -1;
            bool valid = false;
            float coef = 0.f, resid = 0.f;
            int M_obj = 0;
            size_t jb = 0;
            if (MODE == 3) {
                obj = a.w.tile_obj[t]; row0 = a.w.tile_row0[t];
                rows_valid = idle ? 0 : min(TILE_M, a.w.st[obj].ph_rows[a.w.phase] - row0);
            } else if (MODE == 5) {
                obj = a.w.tile_obj[t]; row0 = a.w.tile_row0[t];
                M_obj = a.w.pts_off[obj + 1] - a.w.pts_off[obj];
                const int K = a.w.prm.pose_only ? 0 : a.w.st[obj].K;
                rows_valid = idle ? 0 : min(TILE_M, M_obj + K - row0);
                jb = (size_t)a.w.pts_off[obj] + (size_t)a.w.ray_off[obj] * a.w.prm.D;
            } else {
                const int64_t left = a.n - (int64_t)t * TILE_M;
                rows_valid = idle ? 0 : (int)(left < TILE_M ? left : TILE_M);
            }
            if (hc == 0) {
                float x = 0.f, y = 0.f, z = 0.f;
                valid = r < rows_valid;
                if (valid) {
                    if (MODE == 0 || MODE == 4) {
                        const int64_t gr = (int64_t)t * TILE_M + r;
                        x = a.xyz[gr * 3 + 0]; y = a.xyz[gr * 3 + 1]; z = a.xyz[gr * 3 + 2];
                    } else if (MODE == 1 || MODE == 2) {
                        grid_coord(a.i0 + (int64_t)t * TILE_M + r, a.G, MODE - 1, x, y, z);
                    } else if (MODE == 3) {
                        const float4 v = a.w.frows[(size_t)a.w.ray_off[obj] * a.w.prm.D + row0 + r];
                        x = v.x; y = v.y; z = v.z; dest = __float_as_int(v.w);
                    } else {
                        const float4 v = a.w.jrows[jb + row0 + r];
                        x = v.x; y = v.y; z = v.z; coef = v.w;
                        if (row0 + r >= M_obj) resid = a.w.jres[jb + row0 + r];
                    }
                }
                xyzs[r] = make_float4(x, y, z, coef);
                __half xh, xl, yh, yl, zh, zl;
                split_f16(x, xh, xl); split_f16(y, yh, yl); split_f16(z, zh, zl);
                const __half one = __float2half_rn(1.f), zero = __float2half_rn(0.f);
                const __half2 c0a = __halves2half2(one, one), c0b = __halves2half2(xh, yh), c0c = __halves2half2(zh, xl),
                              c0d = __halves2half2(yl, zl);
                const __half2 c1a = __halves2half2(xh, yh), c1b = __halves2half2(zh, zero), c1z = __halves2half2(zero, zero);
                const uint32_t ax = sAX + (uint32_t)(r >> 3) * 256u + (uint32_t)rx * 16u;
                st_shared_v4(ax, *reinterpret_cast<const uint32_t*>(&c0a), *reinterpret_cast<const uint32_t*>(&c0b),
                             *reinterpret_cast<const uint32_t*>(&c0c), *reinterpret_cast<const uint32_t*>(&c0d));
                st_shared_v4(ax + 128u, *reinterpret_cast<const uint32_t*>(&c1a), *reinterpret_cast<const uint32_t*>(&c1b),
                             *reinterpret_cast<const uint32_t*>(&c1z), *reinterpret_cast<const uint32_t*>(&c1z));
            }
            signal_ready();                                 // layer 0 may start (A_x written, accumulator free)

            // ---------------- forward ----------------
            float sdf_y = 0.f;
            for (int l = 0; l < 8; ++l) {
                const bool last = (l == 7);
                const int ncols = (l == 3) ? a.dec.h3p : HID;
                mbar_wait(bar_dfull, dph);
                dph ^= 1u;
                tc_fence_after();
                TRACE_E(l);
                float part = 0.f;
#pragma unroll 1
                for (int jp = 0; jp < 4; ++jp) {
                    const int col = hc * 256 + jp * 64;
                    if (col >= ncols) {
                        if (JAC) { mw[l][2 * jp] = 0u; mw[l][2 * jp + 1] = 0u; }
                        continue;
                    }
                    uint32_t v0[32], v1[32];
                    tmem_ld32(trow + col, v0);
                    tmem_ld32(trow + col + 32, v1);
                    tmem_ld_wait();
                    if (JAC) {
                        uint32_t m0 = 0u, m1 = 0u;
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            m0 = __funnelshift_l(v0[i], m0, 1);
                            m1 = __funnelshift_l(v1[i], m1, 1);
                        }
                        mw[l][2 * jp] = ~m0;                 // bit (31-i) set <=> pre-activation i positive
                        mw[l][2 * jp + 1] = ~m1;
                    }
                    if (!last) {
                        uint32_t p[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            p[i] = pack_relu_f16x2(__uint_as_float(v0[2 * i]), __uint_as_float(v0[2 * i + 1]));
                        store32(col, p);
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            p[i] = pack_relu_f16x2(__uint_as_float(v1[2 * i]), __uint_as_float(v1[2 * i + 1]));
                        store32(col + 32, p);
                    } else {                                  // layer 8 (1 x 512) folded into the last epilogue
                        const float4* w4 = reinterpret_cast<const float4*>(w8s + col);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 wa = w4[i], wb = w4[8 + i];
                            part = fmaf(fmaxf(__uint_as_float(v0[4 * i + 0]), 0.f), wa.x, part);
                            part = fmaf(fmaxf(__uint_as_float(v0[4 * i + 1]), 0.f), wa.y, part);
                            part = fmaf(fmaxf(__uint_as_float(v0[4 * i + 2]), 0.f), wa.z, part);
                            part = fmaf(fmaxf(__uint_as_float(v0[4 * i + 3]), 0.f), wa.w, part);
                            part = fmaf(fmaxf(__uint_as_float(v1[4 * i + 0]), 0.f), wb.x, part);
                            part = fmaf(fmaxf(__uint_as_float(v1[4 * i + 1]), 0.f), wb.y, part);
                            part = fmaf(fmaxf(__uint_as_float(v1[4 * i + 2]), 0.f), wb.z, part);
                            part = fmaf(fmaxf(__uint_as_float(v1[4 * i + 3]), 0.f), wb.w, part);
                        }
                    }
                }
                if (!last) {
                    signal_ready();
                    TRACE_E(1000 + l);
                } else {
                    parts[hc * TILE_M + r] = part;
                    tc_fence_before();
                    named_bar_sync(1, EPI_THREADS);
                    sdf_y = tanhf(parts[r] + parts[TILE_M + r] + a.dec.b8);
                }
            }

            if (!JAC) {
                if (hc == 0 && valid) {
                    if (MODE == 3) {
                        a.w.sdf_samp[dest] = sdf_y;
                        // The render term takes decisions at |s| = th (occupancy clamp, with_grad mask: loss.py:98-102).
                        // A sample the FP16-operand pass leaves within eps of a threshold is queued for the second,
                        // FP32-grade evaluation (k_tc_jacx<2>), which overwrites sdf_samp[dest] before the ray scan.
                        if (a.w.refine_eps > 0.f && fabsf(fabsf(sdf_y) - a.w.prm.th) < a.w.refine_eps) {
                            const int slot = atomicAdd(&a.w.st[obj].rf_rows, 1);
                            const float4 pc = xyzs[r];
                            a.w.rrows[(size_t)a.w.ray_off[obj] * a.w.prm.D + slot] =
                                make_float4(pc.x, pc.y, pc.z, __int_as_float(dest));
                        }
                    } else {
                        a.out[(int64_t)t * TILE_M + r] = sdf_y;
                    }
                }
                named_bar_sync(1, EPI_THREADS);             // `parts` reusable
                continue;
            }

            // ---------------- backward (input-gradient sweep) ----------------
            if (JAC) {
                const float dyS = GSCALE * (1.f - sdf_y * sdf_y);
                // g7 = dy * w8 . relu'(a7)
#pragma unroll 1
                for (int j = 0; j < 8; ++j) {
                    const int col = hc * 256 + j * 32;
                    const uint32_t m = mw[7][j];
                    const float4* w4 = reinterpret_cast<const float4*>(w8s + col);
                    uint32_t p[16];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 wv = w4[i];
                        const float e0 = ((m << (4 * i + 0)) & 0x80000000u) ? dyS * wv.x : 0.f;
                        const float e1 = ((m << (4 * i + 1)) & 0x80000000u) ? dyS * wv.y : 0.f;
                        const float e2 = ((m << (4 * i + 2)) & 0x80000000u) ? dyS * wv.z : 0.f;
                        const float e3 = ((m << (4 * i + 3)) & 0x80000000u) ? dyS * wv.w : 0.f;
                        p[2 * i] = pack_f16x2(e0, e1);
                        p[2 * i + 1] = pack_f16x2(e2, e3);
                    }
                    store32(col, p);
                }
                signal_ready();
                TRACE_E(1007);

                // steps l = 7..1: D = g_l W_l  -> mask with relu'(a_{l-1}) -> next A
                for (int l = 7; l >= 1; --l) {
                    mbar_wait(bar_dfull, dph);
                    dph ^= 1u;
                    tc_fence_after();
                    TRACE_E(100 + l);
                    const bool split = (l == 4);             // columns >= h3 are d/d(input) (latent re-injection)
                    const int h3 = a.dec.h3, h3p = a.dec.h3p;
                    // one 32-column group: mask with relu'(a_{l-1}), park the input-gradient columns of layer 4, store
                    auto group = [&](int col, uint32_t (&v)[32], uint32_t m) {
                        if (split && col >= h3p) {           // pure input-gradient columns
#pragma unroll
                            for (int i = 0; i < 32; ++i) PARK(col + i - h3) = __uint_as_float(v[i]);
                            return;
                        }
                        if (split && col + 32 > h3) {        // the group that straddles h3 (445)
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                const bool hid = (col + i) < h3;
                                if (!hid) PARK(col + i - h3) = __uint_as_float(v[i]);
                                if (!hid || !((m << i) & 0x80000000u)) v[i] = 0u;
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (!((m << i) & 0x80000000u)) v[i] = 0u;
                        }
                        uint32_t p[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            p[i] = pack_f16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
                        store32(col, p);
                    };
#pragma unroll 1
                    for (int jp = 0; jp < 4; ++jp) {         // two TMEM loads in flight per wait, like the forward sweep
                        const int col = hc * 256 + jp * 64;
                        uint32_t v0[32], v1[32];
                        tmem_ld32(trow + col, v0);
                        tmem_ld32(trow + col + 32, v1);
                        tmem_ld_wait();
                        group(col, v0, mw[l - 1][2 * jp]);
                        group(col + 32, v1, mw[l - 1][2 * jp + 1]);
                    }
                    signal_ready();
                    TRACE_E(1100 + l);
                }

                // step l = 0: D[:, 0..in_dim) = g_0 W_0 ;  jac = (D + gin) / GSCALE
                mbar_wait(bar_dfull, dph);
                dph ^= 1u;
                tc_fence_after();
                TRACE_E(100);
                const float inv_s = 1.f / GSCALE;
                // The two threads of a row share the <= 96 columns: hc == 0 takes columns 0-31 (when the code is at least
                // 32 long), hc == 1 the rest, including the three xyz columns and (GN) the pose part of the J row.
                {
                    const bool lo_half = (L >= 32);          // (uniform) columns 0-31 are code columns owned by hc == 0
                    float* jr = jrow + r * JLD;
                    const float cf = (MODE == 5) ? xyzs[r].w : 1.f;
                    const int o = (MODE == 5) ? POSE : 0;    // MODE 4 rows are [code | xyz], GN rows [pose 7 | code]
                    if (hc == 0) {
                        if (lo_half) {
                            uint32_t v0[32];
                            tmem_ld32(trow + 0, v0);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 32; ++i) jr[o + i] = cf * (__uint_as_float(v0[i]) + PARK(i)) * inv_s;
                        }
                    } else {
                        uint32_t v0[32], v1[32], v2[32];
                        if (!lo_half) tmem_ld32(trow + 0, v0);
                        tmem_ld32(trow + 32, v1);
                        tmem_ld32(trow + 64, v2);
                        tmem_ld_wait();
                        float gxyz[3] = {0.f, 0.f, 0.f};
                        auto col_out = [&](int i, float d) {      // column i of D: a code column or one of the xyz columns
                            const float g = (d + PARK(i)) * inv_s;
                            if (i < L) jr[o + i] = cf * g;
                            else if (i < L + 3) {
                                if (MODE == 4) jr[i] = g;
                                else gxyz[i - L] = g;
                            }
                        };
                        if (!lo_half) {
#pragma unroll
                            for (int i = 0; i < 32; ++i) if (i < L + 3) col_out(i, __uint_as_float(v0[i]));
                        }
#pragma unroll
                        for (int i = 0; i < 32; ++i) if (32 + i < L + 3) col_out(32 + i, __uint_as_float(v1[i]));
#pragma unroll
                        for (int i = 0; i < 32; ++i) if (64 + i < L + 3) col_out(64 + i, __uint_as_float(v2[i]));
                        if (MODE == 5) {
                            const float4 pc = xyzs[r];
                            build_pose_row(jr, cf * gxyz[0], cf * gxyz[1], cf * gxyz[2], pc.x, pc.y, pc.z);
                            for (int c = POSE + L; c < JLD; ++c) jr[c] = 0.f;    // padding columns of the 4x4 blocks
                        }
                    }
                }
                tc_fence_before();

                if (MODE == 4) {
                    named_bar_sync(1, EPI_THREADS);
                    const int in_dim = L + 3;
                    const int64_t g0 = (int64_t)t * TILE_M;
                    if (hc == 0 && valid) a.out[g0 + r] = sdf_y;
                    for (int e = etid; e < rows_valid * in_dim; e += EPI_THREADS) {
                        const int row = e / in_dim, i = e - row * in_dim;
                        a.jac_out[(g0 + row) * in_dim + i] = jrow[row * JLD + i];
                    }
                    named_bar_sync(1, EPI_THREADS);
                } else {
                    // ---- residual weights, then J^T J / J^T r of the tile (optimizer.py:163-169) ----
                    const GnWork& w = a.w;
                    const ObjState& so = w.st[obj];
                    const int K = w.prm.pose_only ? 0 : so.K;
                    const float w_s = (w.prm.pose_only ? 1.f : w.prm.k2) / (float)so.M_active;
                    const float w_r = w.prm.k1 / (float)K;
                    float ls = 0.f, lr = 0.f;
                    if (hc == 0) {
                        float wr = 0.f, rt = 0.f;
                        if (valid) {
                            const int rr = row0 + r;
                            if (rr < M_obj) {                                 // surface row: residual = SDF itself
                                w.sres[w.pts_off[obj] + rr] = sdf_y;
                                rt = w.prm.pose_only ? sdf_y : huber_res(sdf_y, w.prm.b2);
                                if (coef != 0.f) { wr = w_s; ls = rt * rt; }
                            } else {                                          // render row
                                rt = huber_res(resid, w.prm.b1);
                                wr = w_r;
                                lr = rt * rt;
                            }
                        }
                        wrows[r] = wr;
                        rts[r] = rt;
                    }
                    TRACE_E(1400);
                    named_bar_sync(1, EPI_THREADS);
                    TRACE_E(1500);
                    // H is symmetric: 4x4 register blocks (bi <= bj) over the 128 rows, mirrored on the way out;
                    // the threads left over take J^T r.  ~23 instructions per row and block instead of 3 shared
                    // loads per multiply-add (the element-per-thread version was 27 % of the kernel).
                    const int n = POSE + L;
                    const int nb = (n + 3) >> 2;
                    const int nblk = nb * (nb + 1) / 2;
                    float* part = w.partials + ((size_t)obj * w.jgroups_max + (row0 / GROUP_ROWS)) * PART_STRIDE;
                    if (idle) {
                        // (pairs) trip behind the last tile: nothing to reduce, nothing to store
                    } else if (etid < nblk) {
                        int bi = 0, rem = etid;                  // block (bi, bj), bi <= bj, row-major over the upper triangle
                        while (rem >= nb - bi) { rem -= nb - bi; ++bi; }
                        const int bj = bi + rem;
                        float acc[4][4];
#pragma unroll
                        for (int p = 0; p < 4; ++p)
#pragma unroll
                            for (int q = 0; q < 4; ++q) acc[p][q] = 0.f;
                        const float* ja = jrow + 4 * bi;
                        const float* jb4 = jrow + 4 * bj;
#pragma unroll 4
                        for (int row = 0; row < TILE_M; ++row) {
                            const float wr = wrows[row];
                            const float4 av = *reinterpret_cast<const float4*>(ja + row * JLD);
                            const float4 bv = *reinterpret_cast<const float4*>(jb4 + row * JLD);
                            const float a4[4] = {wr * av.x, wr * av.y, wr * av.z, wr * av.w};
                            const float b4[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
                            for (int p = 0; p < 4; ++p)
#pragma unroll
                                for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(a4[p], b4[q], acc[p][q]);
                        }
                        TRACE_E(1550);
#pragma unroll
                        for (int p = 0; p < 4; ++p)
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const int i = 4 * bi + p, j = 4 * bj + q;
                                if (i < n && j < n) {
                                    part[i * n + j] = acc[p][q];
                                    if (bi != bj) part[j * n + i] = acc[p][q];
                                }
                            }
                    } else {
                        for (int i = etid - nblk; i < n; i += EPI_THREADS - nblk) {
                            float sacc = 0.f;
                            for (int row = 0; row < TILE_M; ++row)
                                sacc = fmaf(-wrows[row] * jrow[row * JLD + i], rts[row], sacc);
                            part[n * n + i] = sacc;
                        }
                    }
                    TRACE_E(1600);
                    if (hc == 0) {                                           // loss sums over the 128 rows
                        for (int o = 16; o > 0; o >>= 1) {
                            ls += __shfl_xor_sync(0xffffffffu, ls, o);
                            lr += __shfl_xor_sync(0xffffffffu, lr, o);
                        }
                        if (lane == 0) { parts[q] = ls; parts[4 + q] = lr; }
                    }
                    named_bar_sync(1, EPI_THREADS);
                    if (etid == 0 && !idle) {
                        part[PART_LOSS] = parts[0] + parts[1] + parts[2] + parts[3];
                        part[PART_LOSS + 1] = parts[4] + parts[5] + parts[6] + parts[7];
                    }
                    named_bar_sync(1, EPI_THREADS);
                    TRACE_E(2000);
                }
            }
        }
    }

    // ---------------- teardown ----------------
#ifdef DSP_TC_TRACE
    if (MODE == 5 && blockIdx.x == 0 && lane == 0) {
        if (warp == 0) trace[4095] = te;
        if (warp == 9) trace[8191] = tm;
    }
#endif
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();                           // neither CTA leaves while the other may still signal it
    if (warp == 9) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------
// k_tc_jacx: the decoder at FP32 grade on the tensor cores - both operands as hi + lo FP16 pairs (~22 significand bits).
//
//   * a tile holds 64 real rows stacked to M = 128: tile row r < 64 is hi(a_r), row r + 64 is lo(a_r) (activations in
//     the forward sweep, input gradients in the backward sweep), so ONE M128 MMA multiplies both halves by a weight stage;
//   * every K chunk is issued twice, against the hi(W) stage and the lo(W) stage (the same stage sequence packed from
//     lo(W) = fp16(W - fp16(W)), LayerDesc::src_lo), accumulating in the same TMEM columns:
//       lanes 0-63 : hi(a) (hi(W) + lo(W))        lanes 64-127 : lo(a) (hi(W) + lo(W))
//   * the epilogue adds the two lane halves (see the epilogue branch), applies ReLU (forward) or its mask (backward),
//     splits into hi / lo again and writes rows r and r + 64 of the next A operand.  The warp that finalises a column
//     keeps its ReLU sign bit for the backward sweep (the columns of D_l = g_l W_l are the units of layer l - 1, so the
//     ownership is the same in both directions).
//   * the extra blocks (bias, latent term, xyz columns; already hi/lo) multiply the hi rows only (zero A_x in the lo rows).
//   XM = 0: rows from an xyz array -> sdf[n], jac[n][L + 3]                          (dspslam_sdf_jac, DSPSLAM_MODE_TC_EXACT)
//   XM = 1: rows from the Jacobian work list (64-row tiles) -> J^T J | J^T r partials (dspslam_gn_*, DSPSLAM_MODE_TC_EXACT)
//   XM = 2: forward sweep only over the ray samples the FP16-operand pass left within refine_eps of a decision
//           threshold of the render term (| |s| - th | < eps, 0.5-1 % of the samples): the second evaluation of
//           DSPSLAM_MODE_TC, written over sdf_samp[dest] before the per-ray scan
// Cost: 4 single-pass rows of tensor work per real row and a heavier epilogue: 247 us per 64-row tile through both
// sweeps against 156 us per 128-row tile of the FP16 kernel.
// ------------------------------------------------------------------------------------------------
struct XCfg {
    static constexpr int NST = 4;
    static constexpr int ROWS = REFINE_ROWS;                        // real rows per tile
    static constexpr int O_AX = A_BYTES;
    static constexpr int O_ST = O_AX + AX_BYTES;
    static constexpr int XB_BYTES = 2 * (2 * 2048) + 2 * (2 * 4096);   // slice buffers of the four warp pairs
    static constexpr int O_XB = O_ST + NST * STAGE_BYTES;
    static constexpr int O_W8 = O_XB + XB_BYTES;
    static constexpr int O_XYZ = O_W8 + 2048;                       // float4 x 64
    static constexpr int O_PART = O_XYZ + 1024;                     // 4 x 64 floats
    static constexpr int O_WROW = O_PART + 1024;                    // 64 floats
    static constexpr int O_RT = O_WROW + 256;                       // 64 floats
    static constexpr int O_BAR = O_RT + 256;
    static constexpr int SMEM = O_BAR + 256 + 1024;
};
static_assert(XCfg::ROWS * JLD * 4 <= A_BYTES, "Jacobian rows must fit in the activation tile");

// timing experiments (wrong numerics on purpose; tools/bench_exact.py): -DDSP_X_EXP=2 no activation stores,
// 3 no lo(W) pass
#ifndef DSP_X_EXP
#define DSP_X_EXP 0
#endif
constexpr int XREP = (DSP_X_EXP == 3) ? 1 : 2;

template <int XM>   // 0: xyz array -> sdf + jac;  1: GN Jacobian rows -> J^T J partials;  2: GN second evaluation (forward only)
__global__ void __launch_bounds__(NT, 1) k_tc_jacx(const __grid_constant__ TcArgs a) {
    using C = XCfg;
    constexpr bool GN = (XM == 1), FWD = (XM == 2);
    constexpr int NPROGX = FWD ? 1 : 2;
    constexpr int NST = C::NST;
    constexpr int ROWS = C::ROWS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (sbase - raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sA = sbase, sAX = sbase + C::O_AX, sST = sbase + C::O_ST;
    float* w8s = reinterpret_cast<float*>(sm + C::O_W8);
    float4* xyzs = reinterpret_cast<float4*>(sm + C::O_XYZ);
    float* parts = reinterpret_cast<float*>(sm + C::O_PART);
    float* wrows = reinterpret_cast<float*>(sm + C::O_WROW);
    float* rts = reinterpret_cast<float*>(sm + C::O_RT);
    float* jrow = reinterpret_cast<float*>(sm);
    const uint32_t bar0 = sbase + C::O_BAR;
    auto bar_full = [&](int s) { return bar0 + 8u * s; };
    auto bar_empty = [&](int s) { return bar0 + 8u * (NST + s); };
    const uint32_t bar_aready = bar0 + 8u * (2 * NST), bar_dfull = bar0 + 8u * (2 * NST + 1);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + C::O_BAR + 8 * (2 * NST + 2));

    if (tid == 0) {
        for (int s = 0; s < NST; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), 1); }
        mbar_init(bar_aready, 8);
        mbar_init(bar_dfull, 1);
        for (int i = 0; i < 16; ++i) mbar_init(bar0 + 128u + 8u * i, 1);     // slice buffers: full / empty per pair
        fence_mbar_init();
    }
    if (warp == 9) {
        tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
        tmem_relinquish();
    }
    for (int i = tid; i < HID; i += NT) w8s[i] = __ldg(a.dec.w8 + i);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const int n_tiles = FWD ? a.w.n_tiles[2] : (GN ? a.w.n_tiles[1] : a.n_tiles);

    if (warp == 8) {
        // ===================== weight producer: both programs, every K chunk as a hi(W) and a lo(W) stage =====================
        uint32_t s = 0, ph = 0;
        for (int tb = blockIdx.x; tb < n_tiles; tb += gridDim.x) {
            const int obj = (GN || FWD) ? a.w.tile_obj[tb] : 0;
            const uint8_t* xo = a.xtra + (size_t)obj * (2 * XTRA_BYTES);
            for (int p = 0; p < NPROGX; ++p) {
#pragma unroll 1
                for (int li = 0; li < 8; ++li) {
                    const LayerDesc ld = *reinterpret_cast<const LayerDesc*>(&a.dec.tc_layers[p][li]);
                    if (ld.xkind) {
                        mbar_wait(bar_empty(s), ph ^ 1u);
                        if (elect_one()) {
                            const uint32_t bytes = (uint32_t)ld.n * 32u;
                            mbar_expect_tx(bar_full(s), bytes);
                            bulk_g2s(sST + s * STAGE_BYTES, (ld.xkind == 2 ? xo : a.blob) + ld.xsrc, bytes, bar_full(s));
                        }
                        __syncwarp();
                        if (++s == NST) { s = 0; ph ^= 1u; }
                    }
                    const uint8_t* src_hi = a.blob + ld.src;
                    const uint8_t* src_lo = a.blob + ld.src_lo;
                    for (int n0 = 0; n0 < ld.n; n0 += 256) {
                        const uint32_t bytes = (uint32_t)min(256, (int)ld.n - n0) * 64u;
#pragma unroll 1
                        for (int k2 = 0; k2 < XREP * ld.k32; ++k2) {
                            mbar_wait(bar_empty(s), ph ^ 1u);
                            if (elect_one()) {
                                mbar_expect_tx(bar_full(s), bytes);
                                bulk_g2s(sST + s * STAGE_BYTES, (XREP == 2 && (k2 & 1)) ? src_lo : src_hi, bytes, bar_full(s));
                            }
                            __syncwarp();
                            if (XREP == 1 || (k2 & 1)) { src_hi += bytes; src_lo += bytes; }
                            if (++s == NST) { s = 0; ph ^= 1u; }
                        }
                    }
                }
            }
        }
    } else if (warp == 9) {
        // ===================== MMA issuer =====================
        uint32_t s = 0, ph = 0, aph = 0;
        const uint64_t adesc0 = smem_desc(sA, 16, 1024, 2);
        const uint64_t bdesc0 = smem_desc(sST, 16, 512, 4);
        const uint64_t axdesc = smem_desc(sAX, 128, 256, 0);
        const uint64_t bxdesc0 = smem_desc(sST, 128, 256, 0);
        for (int tb = blockIdx.x; tb < n_tiles; tb += gridDim.x) {
            for (int p = 0; p < NPROGX; ++p) {
#pragma unroll 1
                for (int li = 0; li < 8; ++li) {
                    const LayerDesc ld = *reinterpret_cast<const LayerDesc*>(&a.dec.tc_layers[p][li]);
                    mbar_wait(bar_aready, aph);
                    aph ^= 1u;
                    if (ld.xkind) {
                        mbar_wait(bar_full(s), ph);
                        tc_fence_after();
                        if (elect_one()) {
                            const uint64_t bx = bxdesc0 + (uint64_t)((s * STAGE_BYTES) >> 4);
                            for (int n0 = 0; n0 < ld.n; n0 += 256) {
                                const int nn = min(256, (int)ld.n - n0);
                                umma_f16(tmem + n0, axdesc, bx + (uint64_t)((n0 * 32) >> 4), idesc_f16(TILE_M, nn), 0u);
                            }
                            umma_commit(bar_empty(s));
                        }
                        __syncwarp();
                        if (++s == NST) { s = 0; ph ^= 1u; }
                    }
                    for (int n0 = 0; n0 < ld.n; n0 += 256) {
                        const uint32_t id = idesc_f16(TILE_M, min(256, (int)ld.n - n0));
                        const uint32_t dcol = tmem + n0;
#pragma unroll 1
                        for (int k2 = 0; k2 < XREP * ld.k32; ++k2) {
                            const int kc = (XREP == 2) ? (k2 >> 1) : k2;
                            mbar_wait(bar_full(s), ph);
                            tc_fence_after();
                            if (elect_one()) {
                                const uint64_t ad = adesc0 + (uint64_t)((kc >> 1) * (A_BLK >> 4) + (kc & 1) * 4);
                                const uint64_t bd = bdesc0 + (uint64_t)((s * STAGE_BYTES) >> 4);
                                umma_f16(dcol, ad, bd, id, (ld.xkind == 0 && k2 == 0) ? 0u : 1u);
                                umma_f16(dcol, ad + 2, bd + 2, id, 1u);
                                umma_commit(bar_empty(s));
                            }
                            __syncwarp();
                            if (++s == NST) { s = 0; ph ^= 1u; }
                        }
                    }
                    if (elect_one()) umma_commit(bar_dfull);
                    __syncwarp();
                }
            }
        }
    } else {
        // ===================== epilogue warps =====================
        // Warps q and q + 2 (same column half hc) hold the hi rows and the lo rows of the same 32 real rows.  One of the
        // two FINALISES the pair's 256 columns - adds the partner's accumulators to its own, applies the ReLU (mask),
        // splits into hi / lo and writes the next A operand - the other only SENDS its accumulators through a
        // double-buffered shared-memory slice.  The finaliser is the hi warp in column half 0 and the lo warp in
        // column half 1, so every warp scheduler has one heavy warp.  A lo-row accumulator is ~2^-11 of the hi-row one:
        // the lo warps send theirs as FP16 pairs (half the traffic, 2^-22 relative), the hi warps send FP32.
        const int q = warp & 3, hc = warp >> 2;
        const bool hi_thr = q < 2;                          // this warp's TMEM lanes hold the hi rows (else the lo rows)
        const bool fin = (hc == 0) ? hi_thr : !hi_thr;      // finaliser of the pair's columns [256 hc, 256 hc + 256)
        const int rr = (q & 1) * 32 + lane;                 // real row of the tile
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
        const int rx = rr & 7;
        const uint32_t arow_hi = sA + (uint32_t)(rr >> 3) * 1024u + (uint32_t)rx * 128u;
        const uint32_t arow_lo = arow_hi + 8u * 1024u;      // tile row rr + 64
        const int col0 = hc * 256;
        const int pair = (q & 1) + 2 * hc;
        // slices: pairs 0, 1 (senders = lo warps, FP16 pairs): 2 x 2 KB each; pairs 2, 3 (senders = hi warps, FP32): 2 x 4 KB
        uint8_t* xb = sm + C::O_XB + (hc == 0 ? pair * 4096 : 8192 + (pair - 2) * 8192);
        const int xb_bytes = (hc == 0) ? 2048 : 4096;
        const uint32_t xbar = bar0 + 128u + 32u * pair;     // full[0], full[1], empty[0], empty[1]
        uint32_t xi = 0, xph = 0;                           // slice buffer in use and its phase (same sequence in both warps)
        const int etid = tid;
        const int L = a.dec.L;
        uint32_t dph = 0;
        uint32_t mw[8][8];                                  // finaliser: ReLU sign bits of the pair's 256 columns, per layer
        // layer-4 input gradients of real row rr (written by the finaliser of column half 1, read back by the one of
        // column half 0 in the last step): per CTA [PARK_LD][TILE_M], column-major
        float* park = a.dec.tc_park + (size_t)blockIdx.x * TILE_M * PARK_LD + rr;

        auto signal_ready = [&]() {
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_aready);
        };
        // 32 FP32 values of columns [col, col + 32) of real row rr -> hi / lo FP16 halves in tile rows rr and rr + 64
        auto store_split = [&](int col, const float (&f)[32]) {
            if (DSP_X_EXP == 2) return;
            const uint32_t blk_hi = arow_hi + (uint32_t)(col >> 6) * A_BLK, blk_lo = arow_lo + (uint32_t)(col >> 6) * A_BLK;
            const int cc0 = (col & 63) >> 3;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {                 // 8 columns = one 16-byte chunk per half at a time
                uint32_t h[4], l[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float f0 = f[8 * qq + 2 * c], f1 = f[8 * qq + 2 * c + 1];
                    h[c] = pack_f16x2(f0, f1);
                    const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&h[c]));
                    l[c] = pack_f16x2(f0 - hf.x, f1 - hf.y);
                }
                const uint32_t off = (uint32_t)(((cc0 + qq) ^ rx) << 4);
                st_shared_v4(blk_hi + off, h[0], h[1], h[2], h[3]);
                st_shared_v4(blk_lo + off, l[0], l[1], l[2], l[3]);
            }
        };
        auto x_advance = [&]() { if (++xi == 2) { xi = 0; xph ^= 1u; } };
        // sender: this warp's accumulators of columns [col, col + 32) into the next slice buffer
        auto send_slice = [&](int col) {
            uint32_t v[32];
            tmem_ld32(trow + col, v);
            tmem_ld_wait();
            mbar_wait(xbar + 16u + 8u * xi, xph ^ 1u);       // the finaliser has read what this buffer held
            if (hc == 0) {
                uint32_t* dst = reinterpret_cast<uint32_t*>(xb + xi * xb_bytes);
#pragma unroll
                for (int c = 0; c < 16; ++c)
                    dst[c * 32 + lane] = pack_f16x2(__uint_as_float(v[2 * c]), __uint_as_float(v[2 * c + 1]));
            } else {
                uint32_t* dst = reinterpret_cast<uint32_t*>(xb + xi * xb_bytes);
#pragma unroll
                for (int c = 0; c < 32; ++c) dst[c * 32 + lane] = v[c];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(xbar + 8u * xi);
            x_advance();
        };
        // finaliser: own accumulators of the slice + the partner's
        auto recv_slice = [&](int col, float (&f)[32]) {
            uint32_t v[32];
            tmem_ld32(trow + col, v);
            mbar_wait(xbar + 8u * xi, xph);
            tmem_ld_wait();
            if (hc == 0) {
                const uint32_t* src = reinterpret_cast<const uint32_t*>(xb + xi * xb_bytes);
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const uint32_t h2 = src[c * 32 + lane];
                    const float2 p = __half22float2(*reinterpret_cast<const __half2*>(&h2));
                    f[2 * c] = __uint_as_float(v[2 * c]) + p.x;
                    f[2 * c + 1] = __uint_as_float(v[2 * c + 1]) + p.y;
                }
            } else {
                const float* src = reinterpret_cast<const float*>(xb + xi * xb_bytes);
#pragma unroll
                for (int c = 0; c < 32; ++c) f[c] = __uint_as_float(v[c]) + src[c * 32 + lane];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(xbar + 16u + 8u * xi);
            x_advance();
        };

        for (int tb = blockIdx.x; tb < n_tiles; tb += gridDim.x) {
            // ---------------- tile rows ----------------
            int obj = 0, row0 = 0, rows_valid = 0, M_obj = 0, dest = -1;
            size_t jb = 0;
            if (FWD) {
                obj = a.w.tile_obj[tb]; row0 = a.w.tile_row0[tb];
                rows_valid = min(ROWS, a.w.st[obj].rf_rows - row0);
            } else if (GN) {
                obj = a.w.tile_obj[tb]; row0 = a.w.tile_row0[tb];
                M_obj = a.w.pts_off[obj + 1] - a.w.pts_off[obj];
                const int K = a.w.prm.pose_only ? 0 : a.w.st[obj].K;
                rows_valid = min(ROWS, M_obj + K - row0);
                jb = (size_t)a.w.pts_off[obj] + (size_t)a.w.ray_off[obj] * a.w.prm.D;
            } else {
                const int64_t left = a.n - (int64_t)tb * ROWS;
                rows_valid = (int)(left < ROWS ? left : ROWS);
            }
            const bool valid = rr < rows_valid;
            float coef = 0.f, resid = 0.f;
            if (hc == 0) {
                const int trw = hi_thr ? rr : rr + 64;
                const uint32_t ax = sAX + (uint32_t)(trw >> 3) * 256u + (uint32_t)rx * 16u;
                if (hi_thr) {
                    float x = 0.f, y = 0.f, z = 0.f;
                    if (valid) {
                        if (FWD) {
                            const float4 v = a.w.rrows[(size_t)a.w.ray_off[obj] * a.w.prm.D + row0 + rr];
                            x = v.x; y = v.y; z = v.z; dest = __float_as_int(v.w);
                        } else if (GN) {
                            const float4 v = a.w.jrows[jb + row0 + rr];
                            x = v.x; y = v.y; z = v.z; coef = v.w;
                            if (row0 + rr >= M_obj) resid = a.w.jres[jb + row0 + rr];
                        } else {
                            const int64_t gr = (int64_t)tb * ROWS + rr;
                            x = a.xyz[gr * 3 + 0]; y = a.xyz[gr * 3 + 1]; z = a.xyz[gr * 3 + 2];
                        }
                    }
                    xyzs[rr] = make_float4(x, y, z, coef);
                    __half xh, xl, yh, yl, zh, zl;
                    split_f16(x, xh, xl); split_f16(y, yh, yl); split_f16(z, zh, zl);
                    const __half one = __float2half_rn(1.f), zero = __float2half_rn(0.f);
                    const __half2 c0a = __halves2half2(one, one), c0b = __halves2half2(xh, yh), c0c = __halves2half2(zh, xl),
                                  c0d = __halves2half2(yl, zl);
                    const __half2 c1a = __halves2half2(xh, yh), c1b = __halves2half2(zh, zero), c1z = __halves2half2(zero, zero);
                    st_shared_v4(ax, *reinterpret_cast<const uint32_t*>(&c0a), *reinterpret_cast<const uint32_t*>(&c0b),
                                 *reinterpret_cast<const uint32_t*>(&c0c), *reinterpret_cast<const uint32_t*>(&c0d));
                    st_shared_v4(ax + 128u, *reinterpret_cast<const uint32_t*>(&c1a), *reinterpret_cast<const uint32_t*>(&c1b),
                                 *reinterpret_cast<const uint32_t*>(&c1z), *reinterpret_cast<const uint32_t*>(&c1z));
                } else {
                    st_shared_v4(ax, 0u, 0u, 0u, 0u);
                    st_shared_v4(ax + 128u, 0u, 0u, 0u, 0u);
                }
            }
            signal_ready();

            // ---------------- forward sweep ----------------
            float sdf_y = 0.f;
            for (int l = 0; l < 8; ++l) {
                const bool last = (l == 7);
                const int ncols = (l == 3) ? a.dec.h3p : HID;
                mbar_wait(bar_dfull, dph);
                dph ^= 1u;
                tc_fence_after();
                float part = 0.f;
                const int nsl = min(8, max(0, (ncols - col0) >> 5));       // 32-column slices of this pair in the layer
#pragma unroll 1
                for (int i = 0; i < nsl; ++i) {
                    const int ck = col0 + 32 * i;
                    if (!fin) { send_slice(ck); continue; }
                    float f[32];
                    recv_slice(ck, f);
                    uint32_t m = 0u;
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        m = (m << 1) | (f[c] > 0.f ? 1u : 0u);       // bit (31 - c) set <=> pre-activation c positive
                        f[c] = fmaxf(f[c], 0.f);
                    }
                    if (!FWD) mw[l][i] = m;
                    if (!last) {
                        store_split(ck, f);
                    } else {                                  // layer 8 (1 x 512) folded into the last epilogue
                        const float4* w4 = reinterpret_cast<const float4*>(w8s + ck);
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const float4 wv = w4[c];
                            part = fmaf(f[4 * c + 0], wv.x, part);
                            part = fmaf(f[4 * c + 1], wv.y, part);
                            part = fmaf(f[4 * c + 2], wv.z, part);
                            part = fmaf(f[4 * c + 3], wv.w, part);
                        }
                    }
                }
                if (!last) {
                    signal_ready();
                } else {
                    if (fin) parts[hc * ROWS + rr] = part;
                    tc_fence_before();
                    named_bar_sync(1, EPI_THREADS);
                    sdf_y = tanhf(parts[rr] + parts[ROWS + rr] + a.dec.b8);
                }
            }

            if (FWD) {                                        // second evaluation: the SDF is all that is wanted
                if (hc == 0 && hi_thr && valid) a.w.sdf_samp[dest] = sdf_y;
                named_bar_sync(1, EPI_THREADS);
                continue;
            }

            // ---------------- backward (input-gradient) sweep ----------------
            if (!FWD) {
                // g7 = dy * w8 . relu'(a7) for this thread's 128 columns (no accumulator involved)
                const float dyS = GSCALE * (1.f - sdf_y * sdf_y);
#pragma unroll 1
                for (int i = 0; i < 8 && fin; ++i) {
                    const int ck = col0 + 32 * i;
                    const uint32_t m = mw[7][i];
                    const float4* w4 = reinterpret_cast<const float4*>(w8s + ck);
                    float f[32];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float4 wv = w4[c];
                        f[4 * c + 0] = ((m << (4 * c + 0)) & 0x80000000u) ? dyS * wv.x : 0.f;
                        f[4 * c + 1] = ((m << (4 * c + 1)) & 0x80000000u) ? dyS * wv.y : 0.f;
                        f[4 * c + 2] = ((m << (4 * c + 2)) & 0x80000000u) ? dyS * wv.z : 0.f;
                        f[4 * c + 3] = ((m << (4 * c + 3)) & 0x80000000u) ? dyS * wv.w : 0.f;
                    }
                    store_split(ck, f);
                }
                signal_ready();

                // steps l = 7..1: D = g_l W_l (hi + lo rows) -> sum -> mask with relu'(a_{l-1}) -> next A (hi / lo rows)
                const int h3 = a.dec.h3, h3p = a.dec.h3p;
                for (int l = 7; l >= 1; --l) {
                    mbar_wait(bar_dfull, dph);
                    dph ^= 1u;
                    tc_fence_after();
                    const bool split = (l == 4);             // columns >= h3 are d/d(input) (latent re-injection)
#pragma unroll 1
                    for (int i = 0; i < 8; ++i) {
                        const int ck = col0 + 32 * i;
                        if (!fin) { send_slice(ck); continue; }
                        float f[32];
                        recv_slice(ck, f);
                        const uint32_t m = mw[l - 1][i];
                        if (split && ck >= h3p) {            // pure input-gradient columns: parked, not propagated
#pragma unroll
                            for (int c = 0; c < 32; ++c) park[(ck + c - h3) * TILE_M] = f[c];
                            continue;
                        }
                        if (split && ck + 32 > h3) {         // the slice that straddles h3
#pragma unroll
                            for (int c = 0; c < 32; ++c) {
                                const bool hid = (ck + c) < h3;
                                if (!hid) park[(ck + c - h3) * TILE_M] = f[c];
                                if (!hid || !((m << c) & 0x80000000u)) f[c] = 0.f;
                            }
                        } else {
#pragma unroll
                            for (int c = 0; c < 32; ++c)
                                if (!((m << c) & 0x80000000u)) f[c] = 0.f;
                        }
                        store_split(ck, f);
                    }
                    signal_ready();
                }

                // last step: D[:, 0 .. in_dim) = g_0 W_0; jac = (D + parked layer-4 part) / GSCALE.  The columns lie in
                // column half 0 (lo warps send, hi warps finalise); column half 1 has nothing to do.
                mbar_wait(bar_dfull, dph);
                dph ^= 1u;
                tc_fence_after();
                const float inv_s = 1.f / GSCALE;
                const int in_dim = L + 3;
                if (hc == 0) {
                    float* jr = jrow + rr * JLD;
                    const float cf = GN ? xyzs[rr].w : 1.f;
                    const int o = GN ? POSE : 0;             // stand-alone rows are [code | xyz], GN rows [pose 7 | code]
                    float gxyz[3] = {0.f, 0.f, 0.f};
#pragma unroll 1
                    for (int i = 0; i < 3; ++i) {
                        if (!fin) { send_slice(32 * i); continue; }
                        float f[32];
                        recv_slice(32 * i, f);
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            const int col = 32 * i + c;
                            if (col < in_dim) {
                                const float g = (f[c] + park[col * TILE_M]) * inv_s;
                                if (col < L) jr[o + col] = cf * g;
                                else if (GN) gxyz[col - L] = g;
                                else jr[col] = g;
                            }
                        }
                    }
                    if (GN && fin) {
                        const float4 pc = xyzs[rr];
                        build_pose_row(jr, cf * gxyz[0], cf * gxyz[1], cf * gxyz[2], pc.x, pc.y, pc.z);
                        for (int c = POSE + L; c < JLD; ++c) jr[c] = 0.f;
                    }
                }
                tc_fence_before();
            }

            if (XM == 0) {
                named_bar_sync(1, EPI_THREADS);
                const int in_dim = L + 3;
                const int64_t g0 = (int64_t)tb * ROWS;
                if (hc == 0 && hi_thr && valid) a.out[g0 + rr] = sdf_y;
                for (int e = etid; e < rows_valid * in_dim; e += EPI_THREADS) {
                    const int row = e / in_dim, i = e - row * in_dim;
                    a.jac_out[(g0 + row) * in_dim + i] = jrow[row * JLD + i];
                }
                named_bar_sync(1, EPI_THREADS);
            } else if (GN) {
                // ---- residual weights, then J^T J / J^T r of the tile's 64 rows (optimizer.py:163-169) ----
                const GnWork& w = a.w;
                const ObjState& so = w.st[obj];
                const int K = w.prm.pose_only ? 0 : so.K;
                const float w_s = (w.prm.pose_only ? 1.f : w.prm.k2) / (float)so.M_active;
                const float w_r = w.prm.k1 / (float)K;
                float ls = 0.f, lr = 0.f;
                if (hc == 0 && hi_thr) {
                    float wr = 0.f, rt = 0.f;
                    if (valid) {
                        const int r_obj = row0 + rr;
                        if (r_obj < M_obj) {                                  // surface row: residual = SDF itself
                            w.sres[w.pts_off[obj] + r_obj] = sdf_y;
                            rt = w.prm.pose_only ? sdf_y : huber_res(sdf_y, w.prm.b2);
                            if (coef != 0.f) { wr = w_s; ls = rt * rt; }
                        } else {                                              // render row
                            rt = huber_res(resid, w.prm.b1);
                            wr = w_r;
                            lr = rt * rt;
                        }
                    }
                    wrows[rr] = wr;
                    rts[rr] = rt;
                    for (int o = 16; o > 0; o >>= 1) {
                        ls += __shfl_xor_sync(0xffffffffu, ls, o);
                        lr += __shfl_xor_sync(0xffffffffu, lr, o);
                    }
                    if (lane == 0) { parts[q] = ls; parts[2 + q] = lr; }      // q in {0, 1}
                }
                named_bar_sync(1, EPI_THREADS);
                const int n = POSE + L;
                const int nb = (n + 3) >> 2;
                const int nblk = nb * (nb + 1) / 2;
                float* part = w.partials + ((size_t)obj * w.jgroups_max + (row0 / ROWS)) * PART_STRIDE;
                if (etid < nblk) {
                    int bi = 0, rem = etid;
                    while (rem >= nb - bi) { rem -= nb - bi; ++bi; }
                    const int bj = bi + rem;
                    float acc[4][4];
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int qq = 0; qq < 4; ++qq) acc[p][qq] = 0.f;
                    const float* ja = jrow + 4 * bi;
                    const float* jb4 = jrow + 4 * bj;
#pragma unroll 4
                    for (int row = 0; row < ROWS; ++row) {
                        const float wr = wrows[row];
                        const float4 av = *reinterpret_cast<const float4*>(ja + row * JLD);
                        const float4 bv = *reinterpret_cast<const float4*>(jb4 + row * JLD);
                        const float a4[4] = {wr * av.x, wr * av.y, wr * av.z, wr * av.w};
                        const float b4[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
                        for (int p = 0; p < 4; ++p)
#pragma unroll
                            for (int qq = 0; qq < 4; ++qq) acc[p][qq] = fmaf(a4[p], b4[qq], acc[p][qq]);
                    }
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            const int i = 4 * bi + p, j = 4 * bj + qq;
                            if (i < n && j < n) {
                                part[i * n + j] = acc[p][qq];
                                if (bi != bj) part[j * n + i] = acc[p][qq];
                            }
                        }
                } else {
                    for (int i = etid - nblk; i < n; i += EPI_THREADS - nblk) {
                        float sacc = 0.f;
                        for (int row = 0; row < ROWS; ++row)
                            sacc = fmaf(-wrows[row] * jrow[row * JLD + i], rts[row], sacc);
                        part[n * n + i] = sacc;
                    }
                }
                if (etid == 0) {
                    part[PART_LOSS] = parts[0] + parts[1];
                    part[PART_LOSS + 1] = parts[2] + parts[3];
                }
                named_bar_sync(1, EPI_THREADS);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------
// extra blocks of layers 0 and 4 for the single-code entry points (gn_device.cuh: pack_extra_row; the GN
// drivers build theirs inside k_gn_prep / k_gn_solve)
// ------------------------------------------------------------------------------------------------
__global__ void k_tc_pack_extra(DecoderDev dec, const float* __restrict__ c0, const float* __restrict__ c4,
                                uint8_t* __restrict__ xtra) {
    const int which = blockIdx.y, n = threadIdx.x;
    pack_extra_row(dec, which, n, which ? c4[n] : c0[n], xtra + (size_t)which * XTRA_BYTES);
}

}  // namespace tc

// ------------------------------------------------------------------------------------------------
// host side: weight packing, program tables, launchers
// ------------------------------------------------------------------------------------------------
namespace {

std::string g_tc_err;

struct BlobBuilder {
    std::vector<uint8_t> bytes;

    size_t alloc(size_t n) {
        const size_t at = (bytes.size() + 127) & ~(size_t)127;
        bytes.resize(at + n, 0);
        return at;
    }
    static uint16_t f2h(float v) {
        __half h = __float2half_rn(v);
        uint16_t u;
        memcpy(&u, &h, 2);
        return u;
    }
    static float h2f(uint16_t u) {
        __half h;
        memcpy(&h, &u, 2);
        return __half2float(h);
    }
    // 64-byte-swizzled K-major image of rows [n0, n0+nn) x k in [kc*32, kc*32+32) of the matrix elem(row, k):
    // byte(r, k) = (r/8)*512 + (r%8)*64 + ((k/8) ^ ((r%8)>>1))*16 + (k%8)*2
    template <class F>
    size_t sw64(int n0, int nn, int kc, F elem) {
        const size_t at = alloc((size_t)nn * 64);
        for (int rr = 0; rr < nn; ++rr)
            for (int k = 0; k < 32; ++k) {
                const size_t off = (size_t)(rr >> 3) * 512 + (size_t)(rr & 7) * 64 + (size_t)(((k >> 3) ^ ((rr & 7) >> 1)) << 4) + (size_t)(k & 7) * 2;
                const uint16_t h = f2h(elem(n0 + rr, kc * 32 + k));
                memcpy(&bytes[at + off], &h, 2);
            }
        return at;
    }
};

}  // namespace

const char* tc_last_error() { return g_tc_err.c_str(); }

size_t tc_xtra_bytes_per_object() { return 2 * (size_t)XTRA_BYTES; }

int tc_pack_and_configure(DecoderDev& dec, const float* const* W, const float* const* b, __half** hblob) {
    using namespace tc;
    const int h3 = dec.h3, h3p = dec.h3p, in_dim = dec.in_dim;
    if (h3p % 32 != 0) {          // weight stages are 32 K values wide: code_len 64 gives 448, code_len 32 gives 480
        g_tc_err = "tcgen05 path needs (512 - code_len - 3) rounded up to 16 to be a multiple of 32 (code_len 64, 32, ...)";
        return 1;
    }
    BlobBuilder B;
    LayerDesc layers[2][8];
    memset(layers, 0, sizeof(layers));
    // ---- forward program ----            // layer 0: extras only
    layers[0][0].n = HID; layers[0][0].k32 = 0; layers[0][0].xkind = 2; layers[0][0].xsrc = 0;
    for (int l = 1; l < 8; ++l) {
        const int N = (l == 3) ? h3p : HID, K = (l == 4) ? h3p : HID;
        const int Nreal = (l == 3) ? h3 : HID, Kreal = (l == 4) ? h3 : HID;
        layers[0][l].n = (uint16_t)N; layers[0][l].k32 = (uint8_t)(K / 32);
        if (l == 4) {
            layers[0][l].xkind = 2; layers[0][l].xsrc = XTRA_BYTES;
        } else {                                                                           // static bias block
            const size_t at = B.alloc((size_t)N * 32);
            for (int n = 0; n < Nreal; ++n) {
                const uint16_t hi = BlobBuilder::f2h(b[l][n]);
                const uint16_t lo = BlobBuilder::f2h(b[l][n] - BlobBuilder::h2f(hi));
                const size_t off = at + (size_t)(n >> 3) * 256 + (size_t)(n & 7) * 16;
                memcpy(&B.bytes[off], &hi, 2);
                memcpy(&B.bytes[off + 2], &lo, 2);
            }
            layers[0][l].xkind = 1; layers[0][l].xsrc = (uint32_t)at;
        }
        const float* Wl = W[l];
        for (int n0 = 0; n0 < N; n0 += 256) {
            const int nn = std::min(256, N - n0);
            for (int kb = 0; kb < K / 32; ++kb) {
                const size_t at = B.sw64(n0, nn, kb, [&](int n, int k) -> float {
                    return (n < Nreal && k < Kreal) ? Wl[(size_t)n * HID + k] : 0.f;        // W[l] is [out][512]
                });
                if (n0 == 0 && kb == 0) layers[0][l].src = (uint32_t)at;
            }
        }
    }
    // ---- lo images of the forward program (k_tc_jacx): same stage sequence from lo(W) = fp16(W - fp16(W)) ----
    for (int l = 1; l < 8; ++l) {
        const int N = (l == 3) ? h3p : HID, K = (l == 4) ? h3p : HID;
        const int Nreal = (l == 3) ? h3 : HID, Kreal = (l == 4) ? h3 : HID;
        const float* Wl = W[l];
        for (int n0 = 0; n0 < N; n0 += 256) {
            const int nn = std::min(256, N - n0);
            for (int kb = 0; kb < K / 32; ++kb) {
                const size_t at = B.sw64(n0, nn, kb, [&](int n, int k) -> float {
                    if (!(n < Nreal && k < Kreal)) return 0.f;
                    const float w = Wl[(size_t)n * HID + k];
                    return w - BlobBuilder::h2f(BlobBuilder::f2h(w));
                });
                if (n0 == 0 && kb == 0) layers[0][l].src_lo = (uint32_t)at;
            }
        }
    }
    // ---- backward program: D[:, i] = sum_o g[:, o] W_l[o][i] ----
    for (int l = 7; l >= 1; --l) {
        const int K = (l == 3) ? h3p : HID, Kreal = (l == 3) ? h3 : HID;                   // outputs of layer l
        const float* Wl = W[l];
        layers[1][7 - l].n = HID; layers[1][7 - l].k32 = (uint8_t)(K / 32);
        for (int n0 = 0; n0 < HID; n0 += 256) {
            for (int kb = 0; kb < K / 32; ++kb) {
                const size_t at = B.sw64(n0, 256, kb, [&](int i, int o) -> float {
                    return (o < Kreal) ? Wl[(size_t)o * HID + i] : 0.f;
                });
                if (n0 == 0 && kb == 0) layers[1][7 - l].src = (uint32_t)at;
            }
        }
    }
    // ---- lo images of the backward program ----
    auto lo_of = [](float w) -> float { return w - BlobBuilder::h2f(BlobBuilder::f2h(w)); };
    for (int l = 7; l >= 1; --l) {
        const int K = (l == 3) ? h3p : HID, Kreal = (l == 3) ? h3 : HID;
        const float* Wl = W[l];
        for (int n0 = 0; n0 < HID; n0 += 256) {
            for (int kb = 0; kb < K / 32; ++kb) {
                const size_t at = B.sw64(n0, 256, kb, [&](int i, int o) -> float {
                    return (o < Kreal) ? lo_of(Wl[(size_t)o * HID + i]) : 0.f;
                });
                if (n0 == 0 && kb == 0) layers[1][7 - l].src_lo = (uint32_t)at;
            }
        }
    }
    {
        const int N0 = (in_dim + 15) & ~15;
        for (int kb = 0; kb < HID / 32; ++kb) {
            const size_t at = B.sw64(0, N0, kb, [&](int i, int o) -> float {
                return (i < in_dim) ? lo_of(W[0][(size_t)o * in_dim + i]) : 0.f;
            });
            if (kb == 0) layers[1][7].src_lo = (uint32_t)at;
        }
    }
    {
        const int N0 = (in_dim + 15) & ~15;                                                // 80
        layers[1][7].n = (uint16_t)N0; layers[1][7].k32 = HID / 32;
        for (int kb = 0; kb < HID / 32; ++kb) {
            const size_t at = B.sw64(0, N0, kb, [&](int i, int o) -> float {
                return (i < in_dim) ? W[0][(size_t)o * in_dim + i] : 0.f;
            });
            if (kb == 0) layers[1][7].src = (uint32_t)at;
        }
    }

    // device: [weight images | one extras record for the single-code entry points]
    const size_t wbytes = (B.bytes.size() + 255) & ~(size_t)255;
    const size_t park_bytes = (size_t)2 * PARK_CTAS * TILE_M * PARK_LD * sizeof(float);
    const size_t total = wbytes + 2 * XTRA_BYTES + park_bytes;
    uint8_t* dev = nullptr;
    cudaError_t e = cudaMalloc(&dev, total);
    if (e != cudaSuccess) { g_tc_err = cudaGetErrorString(e); return 2; }
    e = cudaMemcpy(dev, B.bytes.data(), B.bytes.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(dev + wbytes, 0, 2 * XTRA_BYTES);
#define TC_ATTR(MODE)                                                                                          \
    if (e == cudaSuccess)                                                                                      \
        e = cudaFuncSetAttribute(k_tc<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<MODE>::SMEM);
    TC_ATTR(0) TC_ATTR(1) TC_ATTR(2) TC_ATTR(3) TC_ATTR(4) TC_ATTR(5)
#undef TC_ATTR
    static_assert(XCfg::SMEM <= 232448, "shared memory of the hi/lo kernels");
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_tc_jacx<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, XCfg::SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_tc_jacx<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, XCfg::SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_tc_jacx<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, XCfg::SMEM);
    if (e != cudaSuccess) {
        g_tc_err = cudaGetErrorString(e);
        cudaFree(dev);
        return 2;
    }
    *hblob = reinterpret_cast<__half*>(dev);
    dec.tc_blob = dev;
    dec.tc_xtra1 = dev + wbytes;
    dec.tc_park = reinterpret_cast<float*>(dev + wbytes + 2 * XTRA_BYTES);
    memcpy(dec.tc_layers, layers, sizeof(layers));
    return 0;
}

namespace {
template <int MODE>
void launch_tc(const tc::TcArgs& a, int grid, cudaStream_t st) {
    if (!tc::Cfg<MODE>::PAIR) {
        tc::k_tc<MODE><<<grid, tc::NT, tc::Cfg<MODE>::SMEM, st>>>(a);
        return;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)((grid + 1) & ~1), 1, 1);  // whole pairs; the odd CTA of the last pair may idle
    cfg.blockDim = dim3(tc::NT, 1, 1);
    cfg.dynamicSmemBytes = tc::Cfg<MODE>::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, tc::k_tc<MODE>, a);
}
}  // namespace

int tc_launch_sdf_rows(const DecoderDev& dec, int sm_count, const float* xyz, int64_t n, int64_t i0, int G,
                       int src_mode, const float* c0, const float* c4, float* out, cudaStream_t st) {
    using namespace tc;
    k_tc_pack_extra<<<dim3(1, 2), HID, 0, st>>>(dec, c0, c4, dec.tc_xtra1);
    TcArgs a{};
    a.dec = dec; a.blob = dec.tc_blob; a.xtra = dec.tc_xtra1;
    a.xyz = xyz; a.n = n; a.i0 = i0; a.G = G; a.out = out;
    a.n_tiles = (int)((n + TILE_M - 1) / TILE_M);
    const int grid = std::min(a.n_tiles, sm_count);
    if (src_mode == 0) launch_tc<0>(a, grid, st);
    else if (src_mode == 1) launch_tc<1>(a, grid, st);
    else launch_tc<2>(a, grid, st);
    return 2;
}

int tc_launch_sdf_jac(const DecoderDev& dec, int sm_count, const float* xyz, int64_t n, const float* c0,
                      const float* c4, float* sdf_out, float* jac_out, int wsplit, cudaStream_t st) {
    using namespace tc;
    k_tc_pack_extra<<<dim3(1, 2), HID, 0, st>>>(dec, c0, c4, dec.tc_xtra1);
    TcArgs a{};
    a.dec = dec; a.blob = dec.tc_blob; a.xtra = dec.tc_xtra1;
    a.xyz = xyz; a.n = n; a.out = sdf_out; a.jac_out = jac_out; a.wsplit = wsplit;
    a.n_tiles = (int)((n + TILE_M - 1) / TILE_M);
    launch_tc<4>(a, std::min(std::min(a.n_tiles, sm_count), PARK_CTAS), st);
    return 2;
}

// The GN launchers: tile lists (n_tiles, tile_obj, tile_row0) and extras images are already in place, written by
// the glue kernels in front (kernels_simt.cu: k_gn_sample / k_gn_rays / k_gn_prep / k_gn_solve).
int tc_launch_gn_fwd(const GnWork& w, int sm_count, cudaStream_t st) {
    using namespace tc;
    TcArgs a{};
    a.dec = w.dec; a.blob = w.dec.tc_blob; a.xtra = w.tc_xtra; a.w = w;
    launch_tc<3>(a, sm_count, st);
    return 1;
}

int tc_launch_sdf_jac_exact(const DecoderDev& dec, int sm_count, const float* xyz, int64_t n, const float* c0,
                            const float* c4, float* sdf_out, float* jac_out, cudaStream_t st) {
    using namespace tc;
    k_tc_pack_extra<<<dim3(1, 2), HID, 0, st>>>(dec, c0, c4, dec.tc_xtra1);
    TcArgs a{};
    a.dec = dec; a.blob = dec.tc_blob; a.xtra = dec.tc_xtra1;
    a.xyz = xyz; a.n = n; a.out = sdf_out; a.jac_out = jac_out;
    a.n_tiles = (int)((n + XCfg::ROWS - 1) / XCfg::ROWS);
    k_tc_jacx<0><<<std::min(std::min(a.n_tiles, sm_count), PARK_CTAS), NT, XCfg::SMEM, st>>>(a);
    return 2;
}

int tc_launch_gn_jac_exact(const GnWork& w, int sm_count, cudaStream_t st) {
    using namespace tc;
    TcArgs a{};
    a.dec = w.dec; a.blob = w.dec.tc_blob; a.xtra = w.tc_xtra; a.w = w;
    k_tc_jacx<1><<<std::min(sm_count, PARK_CTAS), NT, XCfg::SMEM, st>>>(a);
    return 1;
}

int tc_launch_gn_refine(const GnWork& w, int sm_count, cudaStream_t st) {
    using namespace tc;
    TcArgs a{};
    a.dec = w.dec; a.blob = w.dec.tc_blob; a.xtra = w.tc_xtra; a.w = w;
    k_tc_jacx<2><<<sm_count, NT, XCfg::SMEM, st>>>(a);
    return 1;
}

int tc_launch_gn_jac(const GnWork& w, int sm_count, int early, cudaStream_t st) {
    using namespace tc;
    TcArgs a{};
    a.dec = w.dec; a.blob = w.dec.tc_blob; a.xtra = w.tc_xtra; a.w = w; a.wsplit = w.jac_wsplit;
    if (early) {                                   // the early tile list, its own park blocks
        a.w.tile_obj = w.etile_obj; a.w.tile_row0 = w.etile_row0; a.w.n_tiles = w.n_tiles + 2;   // [1] -> n_tiles[3]
        a.park_cta0 = PARK_CTAS;
    }
    launch_tc<5>(a, std::min(sm_count, PARK_CTAS), st);
    return 1;
}

}  // namespace dsp

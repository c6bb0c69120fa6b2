// FP32 (validation-mode) kernels and the mode-independent Gauss-Newton kernels of
// libdspslam_sm100.  See common.cuh / mlp_simt.cuh / hd_math.h for the pieces and
// include/dspslam.h for the reference lines each entry point replaces.
#include "kernels.h"
#include "gn_device.cuh"
#include "mlp_simt.cuh"

namespace dsp {

// ---------------------------------------------------------------------------
// latent bias:  c0 = b0 + W0[:, :L] z   and   c4 = b4 + W4[:, h3:h3+L] z   (one CTA per code)
// ---------------------------------------------------------------------------
__global__ void k_latent_bias(DecoderDev dec, const float* __restrict__ codes, int code_stride,
                              float* __restrict__ c0, float* __restrict__ c4) {
    __shared__ float z[MAX_L];
    const int b = blockIdx.x;
    if (threadIdx.x < dec.L) z[threadIdx.x] = codes[(size_t)b * code_stride + threadIdx.x];
    __syncthreads();
    for (int n = threadIdx.x; n < HID; n += blockDim.x) {
        float s0 = 0.f, s4 = 0.f;
        for (int l = 0; l < dec.L; ++l) {
            s0 = fmaf(__ldg(dec.Wt0 + (size_t)l * HID + n), z[l], s0);
            s4 = fmaf(__ldg(dec.W4in + (size_t)l * HID + n), z[l], s4);
        }
        c0[(size_t)b * HID + n] = s0 + __ldg(dec.bias[0] + n);
        c4[(size_t)b * HID + n] = s4 + __ldg(dec.bias[4] + n);
    }
}

// ---------------------------------------------------------------------------
// forward-only SDF of n rows with one code.  SRC 0: xyz array, SRC 1: voxel grid index
// ---------------------------------------------------------------------------
template <int SRC>
__global__ void __launch_bounds__(simt::NT, 1)
k_sdf_rows_simt(DecoderDev dec, const float* __restrict__ xyz, int64_t n, int64_t i0, int G,
                int coord_mode, const float* __restrict__ c0, const float* __restrict__ c4,
                float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    simt::Smem& sm = *reinterpret_cast<simt::Smem*>(smem_raw);
    const int tid = threadIdx.x;
    const int64_t row0 = (int64_t)blockIdx.x * GROUP_ROWS;
    for (int sub = 0; sub < GROUP_ROWS / simt::TM; ++sub) {
        const int64_t r0 = row0 + sub * simt::TM;
        if (r0 >= n) break;
        if (tid < simt::TM) {
            const int64_t r = r0 + tid;
            float x = 0.f, y = 0.f, z = 0.f;
            if (r < n) {
                if (SRC == 0) {
                    x = xyz[r * 3 + 0]; y = xyz[r * 3 + 1]; z = xyz[r * 3 + 2];
                } else {
                    grid_coord(i0 + r, G, coord_mode, x, y, z);
                }
            }
            sm.xyz0[tid][0] = x; sm.xyz0[tid][1] = y; sm.xyz0[tid][2] = z; sm.xyz0[tid][3] = 0.f;
        }
        __syncthreads();
        simt::forward32<false>(sm, dec, c0, c4);
        if (tid < simt::TM && r0 + tid < n) out[r0 + tid] = sm.sdf[tid];
        __syncthreads();
    }
}

// SDF + d sdf / d [code | xyz] of n rows with one code
__global__ void __launch_bounds__(simt::NT, 1)
k_sdf_jac_simt(DecoderDev dec, const float* __restrict__ xyz, int64_t n, const float* __restrict__ c0,
               const float* __restrict__ c4, float* __restrict__ sdf_out, float* __restrict__ jac_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    simt::Smem& sm = *reinterpret_cast<simt::Smem*>(smem_raw);
    const int tid = threadIdx.x;
    const int64_t row0 = (int64_t)blockIdx.x * GROUP_ROWS;
    const int in_dim = dec.in_dim;
    for (int sub = 0; sub < GROUP_ROWS / simt::TM; ++sub) {
        const int64_t r0 = row0 + sub * simt::TM;
        if (r0 >= n) break;
        if (tid < simt::TM) {
            const int64_t r = r0 + tid;
            float x = 0.f, y = 0.f, z = 0.f;
            if (r < n) { x = xyz[r * 3 + 0]; y = xyz[r * 3 + 1]; z = xyz[r * 3 + 2]; }
            sm.xyz0[tid][0] = x; sm.xyz0[tid][1] = y; sm.xyz0[tid][2] = z; sm.xyz0[tid][3] = 0.f;
        }
        __syncthreads();
        simt::forward32<true>(sm, dec, c0, c4);
        simt::backward32(sm, dec);
        if (tid < simt::TM && r0 + tid < n) sdf_out[r0 + tid] = sm.sdf[tid];
        for (int e = tid; e < simt::TM * in_dim; e += simt::NT) {
            const int row = e / in_dim, i = e - row * in_dim;
            if (r0 + row < n) jac_out[(r0 + row) * in_dim + i] = sm.jac[row][i];
        }
        __syncthreads();
    }
}

__global__ void k_mean(const float* __restrict__ v, int64_t n, float* __restrict__ out) {
    __shared__ double red[32];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += (double)v[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        out[0] = (float)(t / (double)n);
    }
}

// ---------------------------------------------------------------------------
// Gauss-Newton: init / prep / final
// ---------------------------------------------------------------------------
__global__ void k_gn_init(GnWork w, const float* __restrict__ pose_in, const float* __restrict__ code_in,
                          int code_in_stride, const float* __restrict__ scale_in) {
    const int b = blockIdx.x;
    ObjState& s = w.st[b];
    const int M = w.pts_off[b + 1] - w.pts_off[b];
    if (threadIdx.x == 0) {
        float T[16];
        for (int i = 0; i < 16; ++i) T[i] = pose_in[(size_t)b * 16 + i];
        if (w.prm.pose_only) {                       // optimizer.py:54-56
            const float sc = scale_in[b];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) T[i * 4 + j] *= sc;
        }
        inv_affine(T, s.T_oc);
        s.status = 0; s.n_valid = 0; s.K = 0; s.M_active = M; s.rf_rows = 0;
        for (int p = 0; p < MAX_PHASES; ++p) s.ph_rows[p] = 0;
        s.sdf_loss = 0.f; s.render_loss = 0.f; s.loss = 0.f;
    }
    for (int l = threadIdx.x; l < w.dec.L; l += blockDim.x)
        w.code[(size_t)b * MAX_L + l] = code_in[(size_t)b * code_in_stride + l];
    for (int p = threadIdx.x; p < M; p += blockDim.x) w.active[w.pts_off[b] + p] = 1;
    for (int r = w.ray_off[b] + threadIdx.x; r < w.ray_off[b + 1]; r += blockDim.x) w.ray_hit[r] = -1;
}

// once per iteration and object: pose-derived scalars, latent biases (and, in tcgen05 mode, the extra-block
// images built from them), surface rows in the object frame.  Called by every thread of the object's CTA
// (k_gn_prep before the first iteration, k_gn_solve after each update); status must be 0.
__device__ void prep_object(const GnWork& w, int b, int iter) {
    __shared__ float z[MAX_L];
    __shared__ int cnt;
    const int tid = threadIdx.x;
    ObjState& s = w.st[b];
    const int p0 = w.pts_off[b], M = w.pts_off[b + 1] - p0;
    if (tid == 0) {
        inv_affine(s.T_oc, s.T_co);
        float A[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) A[i * 3 + j] = s.T_co[i * 4 + j];
        const float sc = powf(det3(A), 1.0f / 3.0f);          // optimizer.py:123
        s.scale = sc;
        s.dmin = s.T_co[11] - 1.0f * sc;                       // optimizer.py:125
        s.dmax = s.T_co[11] + 1.0f * sc;
        s.rot_res = w.prm.pose_only ? 0.f : rot_prior(s.T_co, sc, s.rotJ);
        s.n_valid = 0;
        for (int p = 0; p < MAX_PHASES; ++p) s.ph_rows[p] = 0;
        s.rf_rows = 0;
        s.K = 0;
        cnt = 0;
    }
    if (tid < w.dec.L) z[tid] = w.code[(size_t)b * MAX_L + tid];
    __syncthreads();
    if (tid < w.prm.D) w.dtab[(size_t)b * 64 + tid] = linspace_at(s.dmin, s.dmax, w.prm.D, tid);
    for (int n = tid; n < HID; n += blockDim.x) {
        float s0 = 0.f, s4 = 0.f;
        for (int l = 0; l < w.dec.L; ++l) {
            s0 = fmaf(__ldg(w.dec.Wt0 + (size_t)l * HID + n), z[l], s0);
            s4 = fmaf(__ldg(w.dec.W4in + (size_t)l * HID + n), z[l], s4);
        }
        s0 += __ldg(w.dec.bias[0] + n);
        s4 += __ldg(w.dec.bias[4] + n);
        w.c0[(size_t)b * HID + n] = s0;
        w.c4[(size_t)b * HID + n] = s4;
        if (w.tc) {
            uint8_t* img = w.tc_xtra + (size_t)b * (2 * XTRA_BYTES);
            pack_extra_row(w.dec, 0, n, s0, img);
            pack_extra_row(w.dec, 1, n, s4, img + XTRA_BYTES);
        }
    }
    // pose-only inlier cut after iteration e == 4 (optimizer.py:77-79)
    const bool cut = w.prm.pose_only && iter == 5;
    const size_t jb = (size_t)p0 + (size_t)w.ray_off[b] * w.prm.D;
    int local = 0;
    for (int p = tid; p < M; p += blockDim.x) {
        const float* q = w.pts + (size_t)(p0 + p) * 3;
        float x, y, zz;
        sample_point(s.T_oc, q[0], q[1], q[2], 1.0f, x, y, zz);
        uint8_t a = w.active[p0 + p];
        if (cut && a) {
            a = (fabsf(w.sres[p0 + p]) <= 0.05f) ? 1 : 0;
            w.active[p0 + p] = a;
        }
        local += a;
        w.jrows[jb + p] = make_float4(x, y, zz, a ? 1.f : 0.f);
    }
    if (cut) {
        atomicAdd(&cnt, local);
        __syncthreads();
        if (tid == 0) s.M_active = cnt;
    }
}

__global__ void k_gn_prep(GnWork w, int iter) {
    const int b = blockIdx.x;
    if (w.st[b].status == 0) prep_object(w, b, iter);       // (CTA-uniform)
    // pose-only runs have no ray kernels: the Jacobian tile list of the first iteration is planned here
    if (w.tc && w.prm.pose_only && last_cta_done(w.done + 1)) plan_tiles(w, 1);
    // the early Jacobian tiles (surface rows only) do not change over the iterations: planned once
    if (w.tc && w.early && iter == 0 && last_cta_done(w.done + 2)) plan_tiles(w, 3);
}

__global__ void k_gn_final(GnWork w, float* __restrict__ pose_out, float* __restrict__ code_out,
                           int code_out_stride, const float* __restrict__ scale_in,
                           int* __restrict__ status_out, float* __restrict__ loss_out) {
    const int b = blockIdx.x;
    const ObjState& s = w.st[b];
    if (threadIdx.x == 0) {
        float T[16];
        inv_affine(s.T_oc, T);
        if (w.prm.pose_only) {                       // optimizer.py:84-85
            const float sc = scale_in[b];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) T[i * 4 + j] /= sc;
        }
        for (int i = 0; i < 16; ++i) pose_out[(size_t)b * 16 + i] = T[i];
        if (status_out) status_out[b] = s.status;
        if (loss_out) loss_out[b] = s.loss;
    }
    if (code_out)
        for (int l = threadIdx.x; l < w.dec.L; l += blockDim.x)
            code_out[(size_t)b * code_out_stride + l] = w.code[(size_t)b * MAX_L + l];
}

// ---------------------------------------------------------------------------
// render term: sample generation (one warp per ray), loss.py:74-84
//
// The reference evaluates every in-ball sample of every ray.  The render function only needs them up
// to the first sample with sdf <= -th: there the occupancy is exactly 1 (0.5f + th/(2 th), loss.py:98,
// loss_utils.py:46-47), the transmittance cumprod(1-o) is exactly 0 from that sample on, so every later
// sample contributes exactly 0 to d_u, to the acc_trans suffix sums and hence to de_do (loss.py:113-139):
// it can never be selected and never changes a selected row.  With n_phases > 1 the samples of a ray are
// therefore evaluated front to back in phases; a ray that has met such a sample is not continued, and the
// samples never evaluated keep the sentinel (occupancy 0).  The result is bit-identical to n_phases = 1.
// Phase 0 ends two samples behind the previous iteration's first hit (or at the middle of the in-ball
// interval when there is no history), the next phases take 4, 8, ... samples and the last one the rest.
// ---------------------------------------------------------------------------
#ifndef DSP_HIT_MARGIN
#define DSP_HIT_MARGIN 3   // phase 0 ends this many samples behind the previous iteration's first closing sample
#endif
__global__ void __launch_bounds__(256) k_gn_sample(GnWork w) {
    __shared__ int s_obj, s_cnt, s_nv, s_base;
    const int tid = threadIdx.x;
    const int warp = (blockIdx.x * blockDim.x + tid) >> 5;
    const int lane = tid & 31;
    const int D = w.prm.D, P = w.n_phases, ph = w.phase;
    int b = -1;
    bool live = false;
    if (warp < w.total_rays) {
        b = find_segment(w.ray_off, w.B, warp);
        live = w.st[b].status == 0;
    }
    if (tid == 0) { s_obj = b; s_cnt = 0; s_nv = 0; }
    float xs[2] = {0.f, 0.f}, ys[2] = {0.f, 0.f}, zs[2] = {0.f, 0.f};
    unsigned tb[2] = {0u, 0u};
    unsigned long long mask = 0ull;
    if (live) {                                            // (warp-uniform)
        const ObjState& s = w.st[b];
        const float rx = w.rays[(size_t)warp * 3 + 0], ry = w.rays[(size_t)warp * 3 + 1],
                    rz = w.rays[(size_t)warp * 3 + 2];
        unsigned ball[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int j = h * 32 + lane;
            bool valid = false;
            if (j < D) {
                const float d = w.dtab[(size_t)b * 64 + j];
                sample_point(s.T_oc, rx, ry, rz, d, xs[h], ys[h], zs[h]);
                const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(xs[h], xs[h]), __fmul_rn(ys[h], ys[h])),
                                                  __fmul_rn(zs[h], zs[h])));
                valid = nrm < 1.0f;                      // loss.py:82
            }
            ball[h] = __ballot_sync(0xffffffffu, valid);
        }
        mask = (unsigned long long)ball[0] | ((unsigned long long)ball[1] << 32);
        int start = 0, end = 0, j_lo = 0;
        if (mask) {
            j_lo = __ffsll((long long)mask) - 1;
            const int j_hi = 64 - __clzll((long long)mask);
            start = j_lo; end = j_hi;
            if (P > 1) {
                const int pred = w.ray_hit[warp];
                int e = (pred < 0) ? j_lo + (j_hi - j_lo + 1) / 2 : min(j_hi, pred + DSP_HIT_MARGIN);
                e = max(e, min(j_hi, j_lo + 1));
                int st = j_lo;
                for (int p = 1; p <= ph; ++p) {
                    st = e;
                    e = (p == P - 1) ? j_hi : min(j_hi, st + (4 << (p - 1)));
                }
                start = st; end = e;
            }
        }
        if (ph > 0 && start < end) {                     // has an evaluated sample already closed the ray?
            bool hit = false;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int j = h * 32 + lane;
                if (j >= j_lo && j < start && ((mask >> j) & 1ull))
                    hit |= w.sdf_samp[(size_t)warp * D + j] <= -(w.prm.th + w.refine_eps);   // surely closed: a sample inside
                                                                   // the refinement window may still open up again
            }
            if (__any_sync(0xffffffffu, hit)) end = start;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int j = h * 32 + lane;
            const bool take = j < D && ((mask >> j) & 1ull) && j >= start && j < end;
            if (ph == 0 && j < D && !take) w.sdf_samp[(size_t)warp * D + j] = SDF_OUTSIDE;
            tb[h] = __ballot_sync(0xffffffffu, take);
        }
    }
    // row slots: one shared-memory atomic per ray, one global atomic per CTA (rays of the CTA's first object)
    __syncthreads();
    const int cnt = __popc(tb[0]) + __popc(tb[1]);
    const int nv = (ph == 0) ? __popcll(mask) : 0;
    const bool mine = live && b == s_obj;
    int base = 0;
    if (lane == 0 && live) {
        if (mine) {
            if (cnt) base = atomicAdd(&s_cnt, cnt);
            if (nv) atomicAdd(&s_nv, nv);
        } else {
            if (cnt) base = atomicAdd(&w.st[b].ph_rows[ph], cnt);
            if (nv) atomicAdd(&w.st[b].n_valid, nv);
        }
    }
    __syncthreads();
    if (tid == 0 && s_obj >= 0) {
        if (s_cnt) s_base = atomicAdd(&w.st[s_obj].ph_rows[ph], s_cnt);
        if (s_nv) atomicAdd(&w.st[s_obj].n_valid, s_nv);
    }
    __syncthreads();
    base = __shfl_sync(0xffffffffu, base, 0);
    if (mine) base += s_base;
    if (cnt) {
        const size_t fbase = (size_t)w.ray_off[b] * D;  // this object's region of frows
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int j = h * 32 + lane;
            if ((tb[h] >> lane) & 1u) {
                const int off = __popc(tb[h] & ((1u << lane) - 1u)) + (h ? __popc(tb[0]) : 0);
                w.frows[fbase + base + off] = make_float4(xs[h], ys[h], zs[h], __int_as_float(warp * D + j));
            }
        }
    }
    if (w.tc && last_cta_done(w.done + 0)) plan_tiles(w, 0);
}

// tile list of the refinement kernel (one CTA; runs behind the last forward phase, whose epilogues queued the rows)
__global__ void __launch_bounds__(256) k_gn_refine_plan(GnWork w) {
    plan_tiles(w, 2);
    if (w.stats && threadIdx.x == 0) {
        unsigned long long n = 0;
        for (int b = 0; b < w.B; ++b) n += (unsigned long long)w.st[b].rf_rows;
        atomicAdd(w.stats + 3, n);
    }
}

// forward-only SDF of the compacted in-ball samples, FP32 path (grid: groups x objects)
__global__ void __launch_bounds__(simt::NT, 1) k_gn_fwd_simt(GnWork w) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    simt::Smem& sm = *reinterpret_cast<simt::Smem*>(smem_raw);
    const int b = blockIdx.y, tid = threadIdx.x;
    const ObjState& s = w.st[b];
    if (s.status != 0) return;
    const int n = s.ph_rows[w.phase];
    const int row0 = blockIdx.x * GROUP_ROWS;
    if (row0 >= n) return;
    const size_t fbase = (size_t)w.ray_off[b] * w.prm.D;
    const float* c0 = w.c0 + (size_t)b * HID;
    const float* c4 = w.c4 + (size_t)b * HID;
    for (int sub = 0; sub < GROUP_ROWS / simt::TM; ++sub) {
        const int r0 = row0 + sub * simt::TM;
        if (r0 >= n) break;
        int dest = -1;
        if (tid < simt::TM) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r0 + tid < n) { v = w.frows[fbase + r0 + tid]; dest = __float_as_int(v.w); }
            sm.xyz0[tid][0] = v.x; sm.xyz0[tid][1] = v.y; sm.xyz0[tid][2] = v.z; sm.xyz0[tid][3] = 0.f;
        }
        __syncthreads();
        simt::forward32<false>(sm, w.dec, c0, c4);
        if (dest >= 0) w.sdf_samp[dest] = sm.sdf[tid];
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------
// render term: per-ray scan, loss.py:98-155.
//   k_gn_ray_count  grid (ray chunks, objects), thread per ray.  The chunk's SDF samples are staged through shared
//                   memory (coalesced loads, conflict-free pitch).  Each thread runs the scan once and stages the
//                   rows it selects at frows[ray*D + m] (the forward phases are done with that buffer), the
//                   ray's residual in ray_res, and the row offset of the ray inside its chunk.  The last CTA of an
//                   object scans the chunk totals (rows end up in ray order: deterministic J^T J summation order),
//                   sets K and the status; the last CTA of the grid plans the Jacobian tile list of the tcgen05
//                   kernel.
//   k_gn_ray_emit   thread per ray: copies the staged rows behind the object's surface rows.
// ---------------------------------------------------------------------------
constexpr int RAY_NT = 128;

__global__ void __launch_bounds__(RAY_NT) k_gn_ray_count(GnWork w) {
    extern __shared__ float s_sdf[];                           // [RAY_NT][D + 1] samples, then [D][RAY_NT] scan scratch
    __shared__ int s_part[RAY_NT];
    __shared__ int s_last_obj;
    const int b = blockIdx.y, tid = threadIdx.x;
    ObjState& s = w.st[b];
    const int D = w.prm.D, pitch = D + 1;
    const int r0 = w.ray_off[b], N = w.ray_off[b + 1] - r0;
    const int i0 = blockIdx.x * RAY_NT;
    const bool live = s.status == 0;                           // (CTA-uniform)
    int my_cnt = 0;
    if (live && i0 < N) {
        const int nr = min(RAY_NT, N - i0);
        const float* src = w.sdf_samp + (size_t)(r0 + i0) * D;
        {   // flat coalesced copy, (row, column) advanced incrementally (no div/mod), eight loads in flight
            const int n_el = nr * D, drow = RAY_NT / D, dcol = RAY_NT % D;
            int row = tid / D, col = tid % D;
#pragma unroll 8
            for (int e = tid; e < n_el; e += RAY_NT) {
                s_sdf[row * pitch + col] = __ldg(src + e);
                row += drow; col += dcol;
                if (col >= D) { col -= D; ++row; }
            }
        }
        __syncthreads();
        if (tid < nr) {
            const int i = i0 + tid, r = r0 + i;
            RayParams p;
            p.th = w.prm.th; p.dmin = s.dmin; p.dmax = s.dmax; p.D = D; p.dtab = w.dtab + (size_t)b * 64;
            const float* sdf = s_sdf + tid * pitch;
            const float dobs = (i < w.n_fg[b]) ? w.depth[r] : 1.1f * s.dmax;   // optimizer.py:128
            int hit = D;                                 // first closing sample: next iteration's phase-0 horizon
            for (int j = D - 1; j >= 0; --j) if (sdf[j] <= -p.th) hit = j;
            w.ray_hit[r] = hit;
            const float rx = w.rays[(size_t)r * 3 + 0], ry = w.rays[(size_t)r * 3 + 1], rz = w.rays[(size_t)r * 3 + 2];
            float4* stage = w.frows + (size_t)r * D;
            float res_ray = 0.f;
            float* acc = s_sdf + RAY_NT * pitch + tid;            // column tid of the [D][RAY_NT] scratch: conflict-free
            const int cnt = ray_scan_scratch(sdf, 1, dobs, p, acc, RAY_NT, [&](int k, float de_ds, float res) {
                float x, y, z;
                sample_point(s.T_oc, rx, ry, rz, p.dtab[k], x, y, z);
                *stage++ = make_float4(x, y, z, de_ds);
                res_ray = res;
            });
            my_cnt = cnt;
            w.ray_cnt[r] = cnt;
            w.ray_res[r] = res_ray;
        }
    }
    // ---- row offsets inside the chunk (ray order) and the chunk total ----
    {
        const int lane = tid & 31, wid = tid >> 5;
        int inc = my_cnt;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_part[wid] = inc;
        __syncthreads();
        int wbase = 0, total = 0;
#pragma unroll
        for (int q = 0; q < RAY_NT / 32; ++q) {
            const int v = s_part[q];
            if (q < wid) wbase += v;
            total += v;
        }
        if (live && i0 + tid < N) w.ray_roff[r0 + i0 + tid] = wbase + inc - my_cnt;
        if (tid == 0) w.chunk_cnt[(size_t)b * w.ray_chunks + blockIdx.x] = live ? total : 0;
    }
    // ---- last CTA of the object: exclusive scan of its chunk totals ----
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned ticket = atomicAdd(w.done_obj + b, 1u);
        s_last_obj = (ticket == gridDim.x - 1);
        if (s_last_obj) w.done_obj[b] = 0;
    }
    __syncthreads();
    if (s_last_obj && live && tid < 32) {
        __threadfence();
        int* cc = w.chunk_cnt + (size_t)b * w.ray_chunks;
        int carry = 0;
        for (int base = 0; base < w.ray_chunks; base += 32) {
            const int x = base + tid;
            const int v = (x < w.ray_chunks) ? __ldcg(cc + x) : 0;
            int inc = v;
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (tid >= o) inc += t;
            }
            if (x < w.ray_chunks) cc[x] = carry + inc - v;     // in place: count -> first row of the chunk
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (tid == 0) {
            s.K = carry;
            if (*(volatile int*)&s.n_valid < 10) s.status = 1;   // loss.py:87-88
            else if (carry == 0) s.status = 2;          // mean over 0 rows = NaN (optimizer.py:150-152)
        }
    }
    if (w.tc && last_cta_done(w.done + 1)) plan_tiles(w, 1);
}

__global__ void k_gn_ray_emit(GnWork w) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= w.total_rays) return;
    const int cnt = w.ray_cnt[r];
    if (cnt == 0) return;                                       // (also rays of failed objects: counts stay 0)
    const int b = find_segment(w.ray_off, w.B, r);
    if (w.st[b].status != 0) return;
    const int D = w.prm.D;
    const int M = w.pts_off[b + 1] - w.pts_off[b];
    const int chunk = (r - w.ray_off[b]) / RAY_NT;
    size_t at = (size_t)w.pts_off[b] + (size_t)w.ray_off[b] * D + M + w.chunk_cnt[(size_t)b * w.ray_chunks + chunk] +
                w.ray_roff[r];
    const float4* stage = w.frows + (size_t)r * D;
    const float res = w.ray_res[r];
    for (int m = 0; m < cnt; ++m, ++at) {
        w.jrows[at] = stage[m];
        w.jres[at] = res;
    }
}

// ---------------------------------------------------------------------------
// SDF + Jacobian rows -> J^T J / J^T r partial sums, FP32 path (grid: groups x objects)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void build_pose_row(float* J, float gx, float gy, float gz, float x, float y, float z) {
    // d x_o / d xi = [I | -hat(x) | x]  (loss_utils.py:176-195)
    J[0] = gx; J[1] = gy; J[2] = gz;
    J[3] = y * gz - z * gy;
    J[4] = z * gx - x * gz;
    J[5] = x * gy - y * gx;
    J[6] = gx * x + gy * y + gz * z;
}

__global__ void __launch_bounds__(simt::NT, 1) k_gn_jac_simt(GnWork w) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    simt::Smem& sm = *reinterpret_cast<simt::Smem*>(smem_raw);
    __shared__ float red[2][8];
    const int b = blockIdx.y, g = blockIdx.x, tid = threadIdx.x;
    const ObjState& s = w.st[b];
    if (s.status != 0) return;
    const int p0 = w.pts_off[b], M = w.pts_off[b + 1] - p0;
    const int K = w.prm.pose_only ? 0 : s.K;
    const int rows = M + K;
    const int row0 = g * GROUP_ROWS;
    if (row0 >= rows) return;
    const size_t jb = (size_t)p0 + (size_t)w.ray_off[b] * w.prm.D;
    const float* c0 = w.c0 + (size_t)b * HID;
    const float* c4 = w.c4 + (size_t)b * HID;
    const int L = w.dec.L, n = POSE + L;
    const float w_s = (w.prm.pose_only ? 1.f : w.prm.k2) / (float)s.M_active;
    const float w_r = w.prm.k1 / (float)K;
    constexpr int NQ = (HMAX * HMAX + HMAX + simt::NT - 1) / simt::NT;   // 20
    float acc[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) acc[q] = 0.f;
    float ls = 0.f, lr = 0.f;
    for (int sub = 0; sub < GROUP_ROWS / simt::TM; ++sub) {
        const int r0 = row0 + sub * simt::TM;
        if (r0 >= rows) break;
        if (tid < simt::TM) {
            const int r = r0 + tid;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            float res = 0.f;
            if (r < rows) {
                v = w.jrows[jb + r];
                if (r >= M) res = w.jres[jb + r];
            }
            sm.xyz0[tid][0] = v.x; sm.xyz0[tid][1] = v.y; sm.xyz0[tid][2] = v.z; sm.xyz0[tid][3] = 0.f;
            sm.coef[tid] = v.w;
            sm.res[tid] = res;
        }
        __syncthreads();
        simt::forward32<true>(sm, w.dec, c0, c4);
        simt::backward32(sm, w.dec);
        // J rows (pose 7 | code L) into sm.gin
        if (tid < simt::TM) {
            const int r = r0 + tid;
            const float coef = sm.coef[tid];
            const float x = sm.xyz0[tid][0], y = sm.xyz0[tid][1], z = sm.xyz0[tid][2];
            build_pose_row(sm.gin[tid], coef * sm.jac[tid][L], coef * sm.jac[tid][L + 1],
                           coef * sm.jac[tid][L + 2], x, y, z);
            float wr = 0.f, rt = 0.f;
            if (r < rows) {
                if (r < M) {                                   // surface row: residual = SDF itself
                    const float sd = sm.sdf[tid];
                    w.sres[p0 + r] = sd;
                    rt = w.prm.pose_only ? sd : huber_res(sd, w.prm.b2);
                    if (coef != 0.f) { wr = w_s; ls += rt * rt; }
                } else {                                       // render row
                    rt = huber_res(sm.res[tid], w.prm.b1);
                    wr = w_r;
                    lr += rt * rt;
                }
            }
            sm.wrow[tid] = wr;
            sm.rt[tid] = rt;
        }
        for (int e = tid; e < simt::TM * L; e += simt::NT) {
            const int row = e / L, j = e - row * L;
            sm.gin[row][POSE + j] = sm.coef[row] * sm.jac[row][j];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int e = tid + q * simt::NT;
            if (e < n * n) {
                const int i = e / n, j = e - i * n;
                float a = acc[q];
                for (int row = 0; row < simt::TM; ++row)
                    a = fmaf(sm.wrow[row] * sm.gin[row][i], sm.gin[row][j], a);
                acc[q] = a;
            } else if (e < n * n + n) {
                const int i = e - n * n;
                float a = acc[q];
                for (int row = 0; row < simt::TM; ++row)
                    a = fmaf(-sm.wrow[row] * sm.gin[row][i], sm.rt[row], a);
                acc[q] = a;
            }
        }
        __syncthreads();
    }
    float* part = w.partials + ((size_t)b * w.jgroups_max + g) * PART_STRIDE;
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const int e = tid + q * simt::NT;
        if (e < n * n + n) part[e] = acc[q];
    }
    // loss sums live in lanes 0..31 of warp 0 only
    if (tid < 32) {
        for (int o = 16; o > 0; o >>= 1) {
            ls += __shfl_xor_sync(0xffffffffu, ls, o);
            lr += __shfl_xor_sync(0xffffffffu, lr, o);
        }
        if (tid == 0) { part[PART_LOSS] = ls; part[PART_LOSS + 1] = lr; }
    }
    (void)red;
}

// ---------------------------------------------------------------------------
// normal equations: reduce partials, priors + damping, solve, update, and the preparation of the next
// iteration.  optimizer.py:157-194 (joint) / :69-75 (pose only).  One CTA per object.
//
// The reference solves with torch.inverse(H) b in FP32.  Here H (symmetric positive definite after damping)
// is factorised in FP64 as L D L^T, right-looking (rank-1 updates of the trailing lower triangle):
//   D_j = w_jj,   u_ij = w_ij / D_j  (= L_ij),   w_ik -= u_ij w_kj   (j < k <= i)
// with w kept in the lower triangle and u in the (otherwise redundant) upper triangle, transposed.  b rides
// along as row `dim` (its u lives in xs), which makes the forward substitution part of the same loop.
// ---------------------------------------------------------------------------
// Sum of the per-tile partials of every live object, element by element in tile order (FP64): what k_gn_solve used to
// do with one CTA per object (45 % of its time waiting for 0.75 MB of loads per object at C2 size) spread over
// REDUCE_SPLIT CTAs per object.  Same additions in the same order: H and b are bit-identical to the fused version.
constexpr int REDUCE_SPLIT = 8;
__global__ void __launch_bounds__(256) k_gn_reduce(GnWork w) {
    const int b = blockIdx.y, tid = threadIdx.x;
    const ObjState& s = w.st[b];
    if (s.status != 0) return;
    const int n = POSE + w.dec.L;
    const int M = w.pts_off[b + 1] - w.pts_off[b];
    const int K = w.prm.pose_only ? 0 : s.K;
    const int ng = (M + K + w.jac_rows - 1) / w.jac_rows;
    const int total = n * n + n, chunk = (total + REDUCE_SPLIT - 1) / REDUCE_SPLIT;
    const int e0 = blockIdx.x * chunk, e1 = min(total, e0 + chunk);
    const float* part = w.partials + (size_t)b * w.jgroups_max * PART_STRIDE;
    double* hred = w.hred + (size_t)b * PART_STRIDE;
    for (int e = e0 + tid; e < e1; e += blockDim.x) {
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int g = 0;
        for (; g + 16 <= ng; g += 16) {                     // 16 loads in flight, same order of additions as below
            float v[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) v[u] = part[(size_t)(g + u) * PART_STRIDE + e];
#pragma unroll
            for (int u = 0; u < 16; u += 4) {
                a0 += (double)v[u]; a1 += (double)v[u + 1]; a2 += (double)v[u + 2]; a3 += (double)v[u + 3];
            }
        }
        for (; g + 4 <= ng; g += 4) {                       // fixed summation order: ((g0+g4+..)+(g1+..))+...
            a0 += (double)part[(size_t)(g + 0) * PART_STRIDE + e];
            a1 += (double)part[(size_t)(g + 1) * PART_STRIDE + e];
            a2 += (double)part[(size_t)(g + 2) * PART_STRIDE + e];
            a3 += (double)part[(size_t)(g + 3) * PART_STRIDE + e];
        }
        for (; g < ng; ++g) a0 += (double)part[(size_t)g * PART_STRIDE + e];
        hred[e] = (a0 + a1) + (a2 + a3);                    // rows 0..n-1: H, row n: b
    }
}

constexpr int SOLVE_NT = 512;

__global__ void __launch_bounds__(SOLVE_NT) k_gn_solve(GnWork w, int iter, int last_iter, float* __restrict__ dbg,
                                                       int dbg_stride) {
    __shared__ double Hs[(HMAX + 1) * HMAX];
    __shared__ double xs[HMAX + 1];
    __shared__ float losses[2];
    __shared__ int fail;
    const int b = blockIdx.x, tid = threadIdx.x;
    ObjState& s = w.st[b];
    float* dbg_o = dbg ? dbg + ((size_t)iter * w.B + b) * dbg_stride : nullptr;
    const int L = w.dec.L, n = POSE + L;
    bool live = s.status == 0;                                 // (CTA-uniform)
    if (!live && dbg_o && tid == 0 && !w.prm.pose_only) {
        dbg_o[PART_LOSS + 0] = (float)s.K; dbg_o[PART_LOSS + 1] = (float)s.n_valid;
        dbg_o[PART_LOSS + 5] = (float)s.status;
    }
    if (live) {
        const int M = w.pts_off[b + 1] - w.pts_off[b];
        const int K = w.prm.pose_only ? 0 : s.K;
        if (w.stats && tid == 0) {
            int evaluated = 0;
            for (int p = 0; p < MAX_PHASES; ++p) evaluated += s.ph_rows[p];
            atomicAdd(w.stats + 0, (unsigned long long)(w.prm.pose_only ? 0 : evaluated));
            atomicAdd(w.stats + 1, (unsigned long long)(s.M_active + K));
            atomicAdd(w.stats + 2, (unsigned long long)(w.prm.pose_only ? 0 : s.n_valid));
        }
        const int ng = (M + K + w.jac_rows - 1) / w.jac_rows;
        const float* part = w.partials + (size_t)b * w.jgroups_max * PART_STRIDE;
        // the tile partials were summed (in tile order, FP64) by k_gn_reduce, spread over eight CTAs per object
        const double* hred = w.hred + (size_t)b * PART_STRIDE;
        for (int e = tid; e < n * n + n; e += SOLVE_NT) Hs[e] = hred[e];
        if (tid < 2) {
            float a = 0.f;
            for (int g = 0; g < ng; ++g) a += part[(size_t)g * PART_STRIDE + PART_LOSS + tid];
            losses[tid] = a / (tid == 0 ? (float)s.M_active : (float)K);     // mean of robust_res^2
        }
        __syncthreads();
        if (tid == 0) {
            const float sl = losses[0], rl = w.prm.pose_only ? 0.f : losses[1];
            fail = (sl != sl) || (rl != rl);                 // math.isnan checks (optimizer.py:137,151)
            if (!fail) { s.sdf_loss = sl; s.render_loss = rl; }
            else { s.status = 2; if (dbg_o && !w.prm.pose_only) dbg_o[PART_LOSS + 5] = 2.f; }
        }
        __syncthreads();
        live = !fail;
    }
    if (live) {
        int dim;
        if (!w.prm.pose_only) {
            dim = n;
            if (tid < L) {                                                  // code prior (optimizer.py:172-174)
                Hs[(POSE + tid) * n + POSE + tid] += (double)w.prm.k3;
                Hs[n * n + POSE + tid] -= (double)w.prm.k3 * (double)w.code[(size_t)b * MAX_L + tid];
            }
            if (tid == 64) {                                                // rotation prior (:177-181)
                for (int i = 0; i < 3; ++i) {
                    for (int j = 0; j < 3; ++j)
                        Hs[(3 + i) * n + 3 + j] += (double)w.prm.k4 * (double)s.rotJ[i] * (double)s.rotJ[j];
                    Hs[n * n + 3 + i] += (double)w.prm.k4 * (double)s.rotJ[i] * (double)s.rot_res;
                }
            }
            __syncthreads();
            if (tid < POSE) Hs[tid * n + tid] += 1.0 + (tid == POSE - 1 ? (double)w.prm.s_damp : 0.0);   // :185-186
        } else {
            dim = 6;
            // move b (row n) next to the 6x6 block as row 6 so that the factorisation below sees [H; b^T]
            if (tid < 6) xs[tid] = Hs[n * n + tid];
            __syncthreads();
            if (tid < 6) { Hs[6 * n + tid] = xs[tid]; Hs[tid * n + tid] += 1e-2; }          // :71
        }
        __syncthreads();
        if (dbg_o) {
            if (!w.prm.pose_only) {
                for (int e = tid; e < n * n + n; e += SOLVE_NT) dbg_o[e] = (float)Hs[e];
                if (tid == 0) {
                    dbg_o[PART_LOSS + 0] = (float)s.K; dbg_o[PART_LOSS + 1] = (float)s.n_valid;
                    dbg_o[PART_LOSS + 2] = s.sdf_loss; dbg_o[PART_LOSS + 3] = s.render_loss;
                    dbg_o[PART_LOSS + 4] = s.scale;    dbg_o[PART_LOSS + 5] = 0.f;
                }
            } else if (tid < 42) {
                dbg_o[tid] = (float)Hs[(tid / 6) * n + (tid % 6)];          // 6x6 H then b
                if (tid == 0) dbg_o[42] = (float)s.M_active;
            }
            __syncthreads();
        }
        // L D L^T, right-looking and blocked in panels of LDL_P columns.  Panel: thread i = row i (rows j0..dim live
        // in warps 0-2), per column  l_i = w_ic / D_c, then the thread updates the rest of ITS row inside the panel;
        // one 96-thread named barrier per column.  Trailing matrix: all threads (warp = row group, lane = column),
        // one rank-LDL_P update  w_ik -= sum_c l_ic w_kc  between two CTA barriers per panel.
        {
            constexpr int LDL_P = 4;
            const int ty = tid >> 5, tx = tid & 31;
            for (int j0 = 0; j0 < dim; j0 += LDL_P) {
                const int j1 = min(j0 + LDL_P, dim);
                if (tid < 96) {
                    for (int c = j0; c < j1; ++c) {
                        if (tid > c && tid <= dim) {
                            const double u = Hs[tid * n + c] / Hs[c * n + c];
                            if (tid < dim) Hs[c * n + tid] = u; else xs[c] = u;
                            for (int c2 = c + 1; c2 < j1 && c2 <= tid; ++c2)
                                Hs[tid * n + c2] = fma(-u, Hs[c2 * n + c], Hs[tid * n + c2]);
                        }
                        asm volatile("bar.sync 2, 96;" ::: "memory");
                    }
                }
                __syncthreads();
                for (int i = j1 + ty; i <= dim; i += SOLVE_NT / 32) {
                    double li[LDL_P];
#pragma unroll
                    for (int c = 0; c < LDL_P; ++c)
                        li[c] = (j0 + c < j1) ? ((i < dim) ? Hs[(j0 + c) * n + i] : xs[j0 + c]) : 0.0;
                    double* hi = Hs + i * n;
                    for (int k = j1 + tx; k <= i && k < dim; k += 32) {
                        const double* wk = Hs + k * n + j0;
                        double acc = hi[k];
#pragma unroll
                        for (int c = 0; c < LDL_P; ++c)
                            if (j0 + c < j1) acc = fma(-li[c], wk[c], acc);
                        hi[k] = acc;
                    }
                }
                __syncthreads();
            }
        }
        // back substitution L^T x = y (y = xs), column-oriented, warp 0: lane holds x[lane], x[lane+32], x[lane+64]
        if (tid < 32) {
            double x0 = (tid < dim) ? xs[tid] : 0.0, x1 = (tid + 32 < dim) ? xs[tid + 32] : 0.0,
                   x2 = (tid + 64 < dim) ? xs[tid + 64] : 0.0;
            for (int c = dim - 1; c > 0; --c) {
                const double own = (c < 32) ? x0 : (c < 64 ? x1 : x2);
                const double xc = __shfl_sync(0xffffffffu, own, c & 31);
                const double* uc = Hs + c;                                   // u_cj = L_cj at Hs[j*n + c]
                if (tid < c) x0 = fma(-uc[tid * n], xc, x0);
                if (tid + 32 < c) x1 = fma(-uc[(tid + 32) * n], xc, x1);
                if (tid + 64 < c) x2 = fma(-uc[(tid + 64) * n], xc, x2);
            }
            if (tid < dim) xs[tid] = x0;
            if (tid + 32 < dim) xs[tid + 32] = x1;
            if (tid + 64 < dim) xs[tid + 64] = x2;
        }
        __syncthreads();
        // update (optimizer.py:189-194 / :73-75)
        if (tid == 0) {
            float dp[7], dT[16], Tn[16];
            if (!w.prm.pose_only) {
                for (int i = 0; i < 7; ++i) dp[i] = w.prm.lr * (float)xs[i];
                exp_sim3(dp, dT);
            } else {
                for (int i = 0; i < 6; ++i) dp[i] = (float)xs[i];
                exp_se3(dp, dT);
            }
            mul_affine(dT, s.T_oc, Tn);
            for (int i = 0; i < 16; ++i) s.T_oc[i] = Tn[i];
            if (!w.prm.pose_only) s.loss = w.prm.k1 * s.render_loss + w.prm.k2 * s.sdf_loss;
            else s.loss = s.sdf_loss;
        }
        if (!w.prm.pose_only && tid < L)
            w.code[(size_t)b * MAX_L + tid] += w.prm.lr * (float)xs[POSE + tid];
        __syncthreads();
        if (!last_iter) prep_object(w, b, iter + 1);            // next iteration's pose scalars, latent biases, rows
    }
    if (w.tc && w.prm.pose_only && !last_iter && last_cta_done(w.done + 1)) plan_tiles(w, 1);
}

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
static size_t simt_smem() { return sizeof(simt::Smem); }

cudaError_t simt_configure() {
    cudaError_t e;
    const int bytes = (int)simt_smem();
    if ((e = cudaFuncSetAttribute(k_sdf_rows_simt<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(k_sdf_rows_simt<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(k_sdf_jac_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(k_gn_fwd_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(k_gn_jac_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(k_gn_ray_count, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(sizeof(float) * RAY_NT * (2 * 64 + 1)))))
        return e;
    return cudaSuccess;
}

void launch_latent_bias(const DecoderDev& dec, const float* codes, int n_codes, int stride, float* c0,
                        float* c4, cudaStream_t st) {
    k_latent_bias<<<n_codes, 256, 0, st>>>(dec, codes, stride, c0, c4);
}

void launch_sdf_rows_simt(const DecoderDev& dec, const float* xyz, int64_t n, const float* c0,
                          const float* c4, float* out, cudaStream_t st) {
    const unsigned grid = (unsigned)((n + GROUP_ROWS - 1) / GROUP_ROWS);
    k_sdf_rows_simt<0><<<grid, simt::NT, simt_smem(), st>>>(dec, xyz, n, 0, 0, 0, c0, c4, out);
}

void launch_grid_simt(const DecoderDev& dec, int G, int64_t i0, int64_t i1, int coord_mode, const float* c0,
                      const float* c4, float* out, cudaStream_t st) {
    const int64_t n = i1 - i0;
    const unsigned grid = (unsigned)((n + GROUP_ROWS - 1) / GROUP_ROWS);
    k_sdf_rows_simt<1><<<grid, simt::NT, simt_smem(), st>>>(dec, nullptr, n, i0, G, coord_mode, c0, c4, out);
}

void launch_sdf_jac_simt(const DecoderDev& dec, const float* xyz, int64_t n, const float* c0, const float* c4,
                         float* sdf_out, float* jac_out, cudaStream_t st) {
    const unsigned grid = (unsigned)((n + GROUP_ROWS - 1) / GROUP_ROWS);
    k_sdf_jac_simt<<<grid, simt::NT, simt_smem(), st>>>(dec, xyz, n, c0, c4, sdf_out, jac_out);
}

void launch_mean(const float* v, int64_t n, float* out, cudaStream_t st) { k_mean<<<1, 1024, 0, st>>>(v, n, out); }

void launch_gn_init(const GnWork& w, const float* pose_in, const float* code_in, int code_stride,
                    const float* scale_in, cudaStream_t st) {
    k_gn_init<<<w.B, 128, 0, st>>>(w, pose_in, code_in, code_stride, scale_in);
}
void launch_gn_prep(const GnWork& w, int iter, cudaStream_t st) { k_gn_prep<<<w.B, 256, 0, st>>>(w, iter); }
void launch_gn_final(const GnWork& w, float* pose_out, float* code_out, int code_stride, const float* scale_in,
                     int* status_out, float* loss_out, cudaStream_t st) {
    k_gn_final<<<w.B, 64, 0, st>>>(w, pose_out, code_out, code_stride, scale_in, status_out, loss_out);
}
void launch_gn_sample(const GnWork& w, cudaStream_t st) {
    const int64_t threads = (int64_t)w.total_rays * 32;       // (no rays: one CTA that only plans an empty tile list)
    k_gn_sample<<<(unsigned)((threads + 255) / 256 > 0 ? (threads + 255) / 256 : 1), 256, 0, st>>>(w);
}
void launch_gn_refine_plan(const GnWork& w, cudaStream_t st) { k_gn_refine_plan<<<1, 256, 0, st>>>(w); }
void launch_gn_fwd_simt(const GnWork& w, cudaStream_t st) {
    if (w.fgroups_max == 0) return;
    k_gn_fwd_simt<<<dim3(w.fgroups_max, w.B), simt::NT, simt_smem(), st>>>(w);
}
void launch_gn_rays(const GnWork& w, int max_rays, cudaStream_t st) {
    const int chunks = w.ray_chunks;
    k_gn_ray_count<<<dim3(chunks, w.B), RAY_NT, sizeof(float) * RAY_NT * (2 * w.prm.D + 1), st>>>(w);
    k_gn_ray_emit<<<(w.total_rays + 255) / 256 > 0 ? (w.total_rays + 255) / 256 : 1, 256, 0, st>>>(w);
}
void launch_gn_jac_simt(const GnWork& w, cudaStream_t st) {
    k_gn_jac_simt<<<dim3(w.jgroups_max, w.B), simt::NT, simt_smem(), st>>>(w);
}
void launch_gn_solve(const GnWork& w, int iter, int last_iter, float* dbg, int dbg_stride, cudaStream_t st) {
    k_gn_reduce<<<dim3(REDUCE_SPLIT, w.B), 256, 0, st>>>(w);
    k_gn_solve<<<w.B, SOLVE_NT, 0, st>>>(w, iter, last_iter, dbg, dbg_stride);
}

}  // namespace dsp

// Detection -> optimiser-input builder: what the reference's sequence loaders do
// in numpy between the detectors and Optimizer.reconstruct_object, as sm_100a kernels that leave the packed
// batch (pts / rays / depth + offset arrays, the layout dspslam_gn_joint reads) in HBM.
//
//   reconstruct/kitti_sequence.py:136-155  LiDAR crop (3 m cube, oriented box), linspace sub-sampling, camera frame
//   reconstruct/kitti_sequence.py:180-209  projection, mask association, background-pixel sampling, rays
//   reconstruct/kitti_sequence.py:70-97    pixels_sampler (same code in mono_sequence.py:51-78)
//   reconstruct/loss_utils.py:23-37        get_rays
//   reconstruct/mono_sequence.py:95-106    largest mask, sampler, cv2.undistortPoints, rays
//
// This is index / byte work bounded by one read of the scan (16 B per LiDAR point) and of the masks (1 B per pixel).
// Results are BIT-IDENTICAL to the reference's numpy: every product and sum below is an explicit round-to-nearest
// intrinsic in numpy's evaluation order (no fused multiply-add), FP32 where numpy computes in FP32 and FP64 where
// its type promotion lands in FP64; compaction is stable (index order), so the linspace sub-sampling picks the
// same points.
#include <cstdint>

#include "frame.h"

namespace dsp {

namespace {

constexpr int FB_THREADS = 256;
constexpr int FB_SUB = 1;                          // sub-rows of FB_THREADS consecutive points per chunk
constexpr int FB_CHUNK = FB_THREADS * FB_SUB;      // points per CTA
constexpr int FB_WARPS = FB_THREADS / 32;

__device__ __forceinline__ float fm(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fa(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double dm(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double da(double a, double b) { return __dadd_rn(a, b); }

// (p[:, None, :] * R).sum(-1) + t  in FP32: numpy's add-reduce over a length-3 axis is (a0 + a1) + a2
__device__ __forceinline__ void xform(const float* R, const float* t, float x, float y, float z, float* o) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
        o[i] = fa(fa(fa(fm(x, R[3 * i]), fm(y, R[3 * i + 1])), fm(z, R[3 * i + 2])), t[i]);
}

// bit d set <=> the point lies in detection d's 3 m cube and inside its (slightly enlarged) oriented box
__device__ __forceinline__ uint32_t box_flags(const float4 p, const FrameBoxes& bx) {
    uint32_t m = 0;
    for (int d = 0; d < bx.n; ++d) {
        const FrameBoxDev& b = bx.b[d];
        if (p.x > b.lo[0] && p.x < b.hi[0] && p.y > b.lo[1] && p.y < b.hi[1] && p.z > b.lo[2] && p.z < b.hi[2]) {
            float o[3];
            xform(b.Rov, b.tov, p.x, p.y, p.z, o);
            if (o[0] > -b.half[0] && o[0] < b.half[0] && o[1] > -b.half[1] && o[1] < b.half[1] &&
                o[2] > -b.half[2] && o[2] < b.half[2])
                m |= 1u << d;
        }
    }
    return m;
}

// slot of the element of rank r in  a[np.linspace(0, N-1, cap).astype(int32)]  (N > cap), or -1 if it is not picked.
// linspace: y[j] = j * ((N-1)/(cap-1)) in FP64, y[cap-1] = N-1 exactly; the step exceeds 1, so ranks repeat never.
__device__ __forceinline__ int linspace_slot(int r, int N, int cap) {
    if (N <= cap) return r;
    if (cap == 1) return r == 0 ? 0 : -1;
    if (r == N - 1) return cap - 1;
    const double step = __ddiv_rn((double)(N - 1), (double)(cap - 1));
    const int j0 = (int)((double)r / step);
    for (int j = j0 - 1; j <= j0 + 2; ++j)
        if (j >= 0 && j <= cap - 2 && __double2int_rz(dm((double)j, step)) == r) return j;
    return -1;
}

// np.linspace(a, b, n).astype(int32)[i]
__device__ __forceinline__ int linspace_at(int a, int b, int n, int i) {
    if (n == 1) return a;
    if (i == n - 1) return b;
    const double step = __ddiv_rn((double)b - (double)a, (double)(n - 1));
    return __double2int_rz(da(dm((double)i, step), (double)a));
}

__device__ __forceinline__ float2 project(const float* K, float x, float y, float z) {
    float h[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) h[i] = fa(fa(fm(x, K[3 * i]), fm(y, K[3 * i + 1])), fm(z, K[3 * i + 2]));
    return make_float2(__fdiv_rn(h[0], h[2]), __fdiv_rn(h[1], h[2]));
}

// get_rays on FP64 pixel coordinates, one FP32 rounding at the end
template <typename T>
__device__ __forceinline__ void ray_from_pixel(const T* invK, double u, double v, float* out) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
        out[i] = __double2float_rn(
            da(da(dm(u, (double)invK[3 * i]), dm(v, (double)invK[3 * i + 1])), dm(1.0, (double)invK[3 * i + 2])));
}

// ---- 1. per-chunk, per-detection counts of LiDAR points inside the boxes ---------------------------------------
__global__ void __launch_bounds__(FB_THREADS)
k_fb_count(const float4* __restrict__ velo, int64_t n, const __grid_constant__ FrameBoxes bx, int d0, int n_chunks,
           int* __restrict__ chunk_cnt) {
    __shared__ int s_cnt[FB_MAXB];
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < FB_MAXB) s_cnt[tid] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * FB_CHUNK;
#pragma unroll
    for (int i = 0; i < FB_SUB; ++i) {
        const int64_t idx = base + i * FB_THREADS + tid;
        const uint32_t m = idx < n ? box_flags(__ldg(velo + idx), bx) : 0u;
        uint32_t dets = __reduce_or_sync(0xffffffffu, m);
        while (dets) {
            const int d = __ffs(dets) - 1;
            dets &= dets - 1;
            const uint32_t b = __ballot_sync(0xffffffffu, (m >> d) & 1u);
            if (lane == 0) atomicAdd(&s_cnt[d], __popc(b));
        }
    }
    __syncthreads();
    if (tid < bx.n) chunk_cnt[(size_t)(d0 + tid) * n_chunks + blockIdx.x] = s_cnt[tid];
}

// ---- 2. exclusive scan of the chunk counts per detection (one warp each), point offsets of the packed output ---
__global__ void __launch_bounds__(1024)
k_fb_scan(const int* __restrict__ chunk_cnt, int* __restrict__ chunk_off, int n_chunks, int B, int cap,
          int* __restrict__ pts_off, int* __restrict__ info) {
    __shared__ int s_tot[FB_MAX_DETS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int d = warp; d < B; d += 32) {
        int run = 0;
        for (int base = 0; base < n_chunks; base += 32) {
            const int c = base + lane;
            const int v = c < n_chunks ? chunk_cnt[(size_t)d * n_chunks + c] : 0;
            int inc = v;
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, s);
                if (lane >= s) inc += t;
            }
            if (c < n_chunks) chunk_off[(size_t)d * n_chunks + c] = run + inc - v;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) {
            s_tot[d] = run;
            info[d * FB_INFO + FBI_IN_BOX] = run;
            info[d * FB_INFO + FBI_M] = run < cap ? run : cap;
        }
    }
    __syncthreads();
    if (tid == 0) {
        int off = 0;
        for (int d = 0; d < B; ++d) {
            pts_off[d] = off;
            off += s_tot[d] < cap ? s_tot[d] : cap;
        }
        pts_off[B] = off;
    }
}

// ---- 3. stable compaction + linspace sub-sampling + camera-frame transform -------------------------------------
__global__ void __launch_bounds__(FB_THREADS)
k_fb_emit(const float4* __restrict__ velo, int64_t n, const __grid_constant__ FrameBoxes bx, int d0, int n_chunks,
          const int* __restrict__ chunk_cnt, const int* __restrict__ chunk_off, const int* __restrict__ pts_off,
          const int* __restrict__ info, const __grid_constant__ FrameParamsDev prm, float* __restrict__ pts) {
    __shared__ int s_cnt[FB_SUB][FB_WARPS][FB_MAXB];
    __shared__ int s_any;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_any = 0;
    for (int i = tid; i < FB_SUB * FB_WARPS * FB_MAXB; i += FB_THREADS) (&s_cnt[0][0][0])[i] = 0;
    __syncthreads();
    if (tid < bx.n && chunk_cnt[(size_t)(d0 + tid) * n_chunks + blockIdx.x] > 0) s_any = 1;
    __syncthreads();
    if (!s_any) return;                       // nothing of this chunk lies in any box (the common case)
    const int64_t base = (int64_t)blockIdx.x * FB_CHUNK;
    float4 p[FB_SUB];
    uint32_t m[FB_SUB];
#pragma unroll
    for (int i = 0; i < FB_SUB; ++i) {
        const int64_t idx = base + i * FB_THREADS + tid;
        p[i] = idx < n ? __ldg(velo + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
        m[i] = idx < n ? box_flags(p[i], bx) : 0u;
        uint32_t dets = __reduce_or_sync(0xffffffffu, m[i]);
        while (dets) {
            const int d = __ffs(dets) - 1;
            dets &= dets - 1;
            const uint32_t b = __ballot_sync(0xffffffffu, (m[i] >> d) & 1u);
            if (lane == 0) s_cnt[i][warp][d] = __popc(b);
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < FB_SUB; ++i) {
        uint32_t dets = __reduce_or_sync(0xffffffffu, m[i]);
        while (dets) {
            const int d = __ffs(dets) - 1;
            dets &= dets - 1;
            const uint32_t b = __ballot_sync(0xffffffffu, (m[i] >> d) & 1u);
            if ((m[i] >> d) & 1u) {
                int r = chunk_off[(size_t)(d0 + d) * n_chunks + blockIdx.x] + __popc(b & ((1u << lane) - 1u));
                for (int q = 0; q < i * FB_WARPS + warp; ++q) r += (&s_cnt[0][0][0])[q * FB_MAXB + d];
                const int N = info[(d0 + d) * FB_INFO + FBI_IN_BOX];
                const int slot = linspace_slot(r, N, prm.max_pts);
                if (slot >= 0) {
                    float o[3];
                    xform(prm.Rcv, prm.tcv, p[i].x, p[i].y, p[i].z, o);
                    float* dst = pts + (size_t)(pts_off[d0 + d] + slot) * 3;
                    dst[0] = o[0]; dst[1] = o[1]; dst[2] = o[2];
                }
            }
        }
    }
}

// ---- 4. mask areas (bytes are 0/1 or 0/255: count the non-zero ones, four per 32-bit load) -----------------------
__global__ void __launch_bounds__(FB_THREADS)
k_fb_mask_area(const uint8_t* __restrict__ masks, int hw, int* __restrict__ area) {
    const uint8_t* mk = masks + (size_t)blockIdx.y * hw;
    const int head = (int)((4 - (reinterpret_cast<uintptr_t>(mk) & 3)) & 3);       // bytes before the first aligned word
    const int n_words = hw > head ? (hw - head) / 4 : 0;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(mk + head);
    int c = 0;
    const int t = blockIdx.x * FB_THREADS + threadIdx.x, stride = gridDim.x * FB_THREADS;
    for (int i = t; i < n_words; i += stride) {
        const uint32_t v = __ldg(w + i);
        // per byte: non-zero -> 1.  (v | v>>4) folds nibbles, then |>>2, |>>1 fold into bit 0 of every byte
        uint32_t f = v | (v >> 4);
        f |= f >> 2;
        f |= f >> 1;
        c += __popc(f & 0x01010101u);
    }
    if (t < head && t < hw) c += mk[t] != 0;
    const int tail0 = head + 4 * n_words;
    if (t < hw - tail0) c += mk[tail0 + t] != 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&area[blockIdx.y], c);
}

// Background pixels of one mask: lattice with stride alpha over the 2-D box grown by 5 pixels, pixels NOT on the
// mask, in row-major order, sub-sampled to at most cap by linspace.  Called by a whole CTA of FB_THREADS threads;
// returns the number of pixels written to out (all threads get the same value).
__device__ int sample_background(const uint8_t* __restrict__ mk, int W, int H, const float* __restrict__ bbox, int alpha,
                                 int cap, int2* out, int* s_warp /* FB_WARPS + 2 ints of shared memory */) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int max_w = W - 1, max_h = H - 1;
    int l = __float2int_rz(bbox[0]), t = __float2int_rz(bbox[1]), r = __float2int_rz(bbox[2]), b = __float2int_rz(bbox[3]);
    l = l > 5 ? l - 5 : 0;
    t = t > 5 ? t - 5 : 0;
    r = r < max_w - 5 ? r + 5 : max_w;
    b = b < max_h - 5 ? b + 5 : max_h;
    const int n_h = (b - t + 1) / alpha, n_w = (r - l + 1) / alpha;
    if (n_h <= 0 || n_w <= 0) return 0;
    const int n_lat = n_h * n_w;
    int total = 0;
    for (int pass = 0; pass < 2; ++pass) {
        int run = 0;
        for (int base = 0; base < n_lat; base += FB_THREADS) {
            const int q = base + tid;
            int u = 0, v = 0;
            bool keep = false;
            if (q < n_lat) {
                v = linspace_at(t, b, n_h, q / n_w);
                u = linspace_at(l, r, n_w, q % n_w);
                keep = (u >= 0 && u < W && v >= 0 && v < H) && mk[(size_t)v * W + u] == 0;
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) s_warp[warp] = __popc(bal);
            __syncthreads();
            int before = 0, all = 0;
#pragma unroll
            for (int w = 0; w < FB_WARPS; ++w) {
                const int c = s_warp[w];
                before += w < warp ? c : 0;
                all += c;
            }
            if (pass == 1 && keep) {
                const int rk = run + before + __popc(bal & ((1u << lane) - 1u));
                const int slot = linspace_slot(rk, total, cap);
                if (slot >= 0) out[slot] = make_int2(u, v);
            }
            run += all;
            __syncthreads();
        }
        total = run;
    }
    return total < cap ? total : cap;
}

// ---- 5. association of each detection with a 2-D mask, background pixels ------------------------------------------
__global__ void __launch_bounds__(FB_THREADS)
k_fb_associate(const float* __restrict__ pts, const int* __restrict__ pts_off, const uint8_t* __restrict__ masks,
               int n_masks, const float* __restrict__ bboxes, const int* __restrict__ area,
               const __grid_constant__ FrameParamsDev prm, const __grid_constant__ FrontBits front,
               int2* __restrict__ bg_px, int* __restrict__ info) {
    __shared__ int s_hits[FB_MAX_MASKS];
    __shared__ int s_nfov, s_mask;
    __shared__ int s_warp[FB_WARPS + 2];
    const int d = blockIdx.x, tid = threadIdx.x;
    int* inf = info + d * FB_INFO;
    const bool is_front = (front.w[d >> 5] >> (d & 31)) & 1u;
    for (int i = tid; i < n_masks; i += FB_THREADS) s_hits[i] = 0;
    if (tid == 0) { s_nfov = 0; s_mask = -1; }
    __syncthreads();
    const int p0 = pts_off[d], M = pts_off[d + 1] - p0;
    const int W = prm.img_w, H = prm.img_h;
    if (is_front && n_masks > 0) {
        for (int j = tid; j < M; j += FB_THREADS) {
            const float* p = pts + (size_t)(p0 + j) * 3;
            const float2 uv = project(prm.K, p[0], p[1], p[2]);
            if (uv.x > 0.f && uv.x < (float)W && uv.y > 0.f && uv.y < (float)H) {
                const int u = __float2int_rz(uv.x), v = __float2int_rz(uv.y);
                atomicAdd(&s_nfov, 1);
                const size_t at = (size_t)v * W + u;
                for (int k = 0; k < n_masks; ++k)
                    if (masks[(size_t)k * W * H + at]) atomicAdd(&s_hits[k], 1);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        int best = -1, hmax = -1;
        for (int k = 0; k < n_masks; ++k)
            if (s_hits[k] > hmax) { hmax = s_hits[k]; best = k; }          // first maximum, as np.argmax
        if (is_front && n_masks > 0 && (double)hmax > (double)s_nfov * 0.5) s_mask = best;
        inf[FBI_NFOV] = s_nfov;
        inf[FBI_MASK] = s_mask;
        inf[FBI_HITS] = hmax < 0 ? 0 : hmax;
    }
    __syncthreads();
    const int mk = s_mask;
    int n_bg = 0, has = 0;
    if (mk >= 0 && area[mk] > prm.min_mask_area) {
        has = 1;
        n_bg = sample_background(masks + (size_t)mk * W * H, W, H, bboxes + 4 * mk, prm.alpha, prm.max_bg,
                                 bg_px + (size_t)d * prm.max_bg, s_warp);
    }
    if (tid == 0) { inf[FBI_HAS_RAYS] = has; inf[FBI_NBG] = n_bg; }
}

// ---- 6. rays + depth in the packed layout of dspslam_gn_joint ---------------------------------------------------
__global__ void __launch_bounds__(FB_THREADS)
k_fb_pack(const float* __restrict__ pts, const int* __restrict__ pts_off, const int2* __restrict__ bg_px,
          const int* __restrict__ info, int B, const __grid_constant__ FrameParamsDev prm, int* __restrict__ ray_off,
          int* __restrict__ n_fg, float* __restrict__ rays, float* __restrict__ depth) {
    __shared__ int s_off;
    const int d = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) {
        int off = 0;
        for (int e = 0; e < d; ++e)
            if (info[e * FB_INFO + FBI_HAS_RAYS]) off += info[e * FB_INFO + FBI_M] + info[e * FB_INFO + FBI_NBG];
        s_off = off;
        ray_off[d] = off;
        const int has = info[d * FB_INFO + FBI_HAS_RAYS];
        n_fg[d] = has ? info[d * FB_INFO + FBI_M] : 0;
        if (d == B - 1) ray_off[B] = off + (has ? info[d * FB_INFO + FBI_M] + info[d * FB_INFO + FBI_NBG] : 0);
    }
    __syncthreads();
    if (!info[d * FB_INFO + FBI_HAS_RAYS]) return;
    const int ro = s_off, p0 = pts_off[d], M = info[d * FB_INFO + FBI_M], n_bg = info[d * FB_INFO + FBI_NBG];
    for (int j = tid; j < M + n_bg; j += FB_THREADS) {
        float r[3];
        if (j < M) {
            const float* p = pts + (size_t)(p0 + j) * 3;
            const float2 uv = project(prm.K, p[0], p[1], p[2]);
            ray_from_pixel(prm.invK, (double)uv.x, (double)uv.y, r);
            depth[ro + j] = p[2];
        } else {
            const int2 px = bg_px[(size_t)d * prm.max_bg + (j - M)];
            ray_from_pixel(prm.invK, (double)px.x, (double)px.y, r);
            depth[ro + j] = 0.f;
        }
        float* dst = rays + (size_t)(ro + j) * 3;
        dst[0] = r[0]; dst[1] = r[1]; dst[2] = r[2];
    }
}

// ---- monocular: largest mask, sampler, undistortion, rays ---------------------------------------------------------
__global__ void __launch_bounds__(FB_THREADS)
k_fb_mono(const uint8_t* __restrict__ masks, int n_masks, const float* __restrict__ bboxes, const int* __restrict__ area,
          const __grid_constant__ FrameParamsDev prm, int2* __restrict__ px_tmp, float* __restrict__ rays,
          int* __restrict__ info) {
    __shared__ int s_mask;
    __shared__ int s_warp[FB_WARPS + 2];
    const int tid = threadIdx.x;
    if (tid == 0) {
        int best = 0, amax = -1;
        for (int k = 0; k < n_masks; ++k)
            if (area[k] > amax) { amax = area[k]; best = k; }
        s_mask = best;
    }
    __syncthreads();
    const int mk = s_mask, W = prm.img_w, H = prm.img_h;
    const int n_bg = sample_background(masks + (size_t)mk * W * H, W, H, bboxes + 4 * mk, prm.alpha, prm.max_bg, px_tmp,
                                       s_warp);
    __syncthreads();
    // cv2.undistortPoints(pix, K, (k1, k2, 0, 0, 0), P=K): five fixed-point iterations in FP64 (OpenCV
    // cvUndistortPointsInternal), FP32 in and out
    const double fx = prm.Kd[0], fy = prm.Kd[4], cx = prm.Kd[2], cy = prm.Kd[5], k1 = prm.k1, k2 = prm.k2;
    for (int j = tid; j < n_bg; j += FB_THREADS) {
        const int2 px = px_tmp[j];
        const double x0 = __ddiv_rn(da((double)px.x, -cx), fx), y0 = __ddiv_rn(da((double)px.y, -cy), fy);
        double x = x0, y = y0;
        for (int it = 0; it < 5; ++it) {
            const double r2 = da(dm(x, x), dm(y, y));
            const double ic = __ddiv_rn(1.0, da(1.0, dm(da(dm(k2, r2), k1), r2)));
            if (ic < 0) { x = x0; y = y0; }
            else { x = dm(x0, ic); y = dm(y0, ic); }
        }
        const float u = __double2float_rn(da(dm(x, fx), cx)), v = __double2float_rn(da(dm(y, fy), cy));
        float r[3];
        ray_from_pixel(prm.invKd, (double)u, (double)v, r);
        rays[3 * j] = r[0]; rays[3 * j + 1] = r[1]; rays[3 * j + 2] = r[2];
    }
    if (tid == 0) { info[0] = mk; info[1] = n_bg; }
}

}  // namespace

namespace {
struct Scratch {
    int* chunk_cnt; int* chunk_off; int* area; int2* bg_px; int n_chunks; size_t bytes;
};
Scratch carve(void* scratch, int64_t n_velo, int B, int n_masks, int max_bg) {
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    Scratch s;
    s.n_chunks = (int)((n_velo + FB_CHUNK - 1) / FB_CHUNK);
    if (s.n_chunks < 1) s.n_chunks = 1;
    if (B < 1) B = 1;
    char* base = reinterpret_cast<char*>(scratch);
    const size_t a = al((size_t)B * s.n_chunks * sizeof(int));
    const size_t m = al((size_t)(n_masks + 1) * sizeof(int));
    s.chunk_cnt = reinterpret_cast<int*>(base);
    s.chunk_off = reinterpret_cast<int*>(base + a);
    s.area = reinterpret_cast<int*>(base + 2 * a);
    s.bg_px = reinterpret_cast<int2*>(base + 2 * a + m);
    s.bytes = 2 * a + m + al((size_t)B * (max_bg > 0 ? max_bg : 1) * sizeof(int2));
    return s;
}
}  // namespace

size_t frame_scratch_bytes(int64_t n_velo, int B, int n_masks, int max_bg) {
    return carve(nullptr, n_velo, B, n_masks, max_bg).bytes;
}

int launch_frame_kitti(const FrameParamsDev& prm, const float* velo, int64_t n_velo, const FrameBoxDev* boxes,
                       const int* is_front, int B, const uint8_t* masks, int n_masks, const float* bboxes,
                       int* pts_off, int* ray_off, int* n_fg, float* pts, float* rays, float* depth, int* info,
                       void* scratch, cudaStream_t st) {
    Scratch s = carve(scratch, n_velo, B, n_masks, prm.max_bg);
    int nl = 0;
    cudaMemsetAsync(s.area, 0, sizeof(int) * (n_masks + 1), st);
    cudaMemsetAsync(s.chunk_cnt, 0, sizeof(int) * (size_t)B * s.n_chunks, st);
    const float4* v4 = reinterpret_cast<const float4*>(velo);
    const int grid = n_velo > 0 ? s.n_chunks : 0;
    FrameBoxes bx;
    for (int d0 = 0; d0 < B && grid; d0 += FB_MAXB) {
        bx.n = B - d0 < FB_MAXB ? B - d0 : FB_MAXB;
        for (int i = 0; i < bx.n; ++i) bx.b[i] = boxes[d0 + i];
        k_fb_count<<<grid, FB_THREADS, 0, st>>>(v4, n_velo, bx, d0, s.n_chunks, s.chunk_cnt); ++nl;
    }
    k_fb_scan<<<1, 1024, 0, st>>>(s.chunk_cnt, s.chunk_off, s.n_chunks, B, prm.max_pts, pts_off, info); ++nl;
    for (int d0 = 0; d0 < B && grid; d0 += FB_MAXB) {
        bx.n = B - d0 < FB_MAXB ? B - d0 : FB_MAXB;
        for (int i = 0; i < bx.n; ++i) bx.b[i] = boxes[d0 + i];
        k_fb_emit<<<grid, FB_THREADS, 0, st>>>(v4, n_velo, bx, d0, s.n_chunks, s.chunk_cnt, s.chunk_off, pts_off, info,
                                               prm, pts); ++nl;
    }
    if (n_masks > 0) {
        const int hw = prm.img_w * prm.img_h;
        int gx = (hw / 4 + FB_THREADS * 4 - 1) / (FB_THREADS * 4);
        if (gx < 1) gx = 1;
        k_fb_mask_area<<<dim3(gx, n_masks), FB_THREADS, 0, st>>>(masks, hw, s.area); ++nl;
    }
    FrontBits fb{};
    for (int d = 0; d < B; ++d)
        if (is_front[d]) fb.w[d >> 5] |= 1u << (d & 31);
    k_fb_associate<<<B, FB_THREADS, 0, st>>>(pts, pts_off, masks, n_masks, bboxes, s.area, prm, fb, s.bg_px, info); ++nl;
    k_fb_pack<<<B, FB_THREADS, 0, st>>>(pts, pts_off, s.bg_px, info, B, prm, ray_off, n_fg, rays, depth); ++nl;
    return nl;
}

int launch_frame_mono(const FrameParamsDev& prm, const uint8_t* masks, int n_masks, const float* bboxes, float* rays,
                      int* info, void* scratch, cudaStream_t st) {
    Scratch s = carve(scratch, 0, 1, n_masks, prm.max_bg);
    const int hw = prm.img_w * prm.img_h;
    cudaMemsetAsync(s.area, 0, sizeof(int) * (n_masks + 1), st);
    int gx = (hw / 4 + FB_THREADS * 4 - 1) / (FB_THREADS * 4);
    if (gx < 1) gx = 1;
    k_fb_mask_area<<<dim3(gx, n_masks), FB_THREADS, 0, st>>>(masks, hw, s.area);
    k_fb_mono<<<1, FB_THREADS, 0, st>>>(masks, n_masks, bboxes, s.area, prm, s.bg_px, rays, info);
    return 2;
}

}  // namespace dsp

// C ABI of libdspslam_sm100 (include/dspslam.h): handle, weight packing, workspace,
// and the launch sequences of the Gauss-Newton drivers.  No exception crosses the ABI.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/dspslam.h"
#include "kernels.h"
#include "kernels_tc.h"
#include "frame.h"
#include "posegraph.h"

using namespace dsp;

static thread_local std::string g_err;

static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(DSPSLAM_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));       \
    } while (0)

struct dspslam_handle {
    int device = 0;
    int sm_count = 0;
    DecoderDev dec{};
    float* wblob = nullptr;      // FP32 packed parameters
    __half* hblob = nullptr;     // FP16 tcgen05 operand tiles
    char* ws = nullptr;          // scratch, grown on demand
    size_t ws_cap = 0;
    float* lat = nullptr;        // c0 | c4 of the single-code entry points (2 x 512) + staging
    int32_t* zeros = nullptr;    // zero offsets for the ray-less pose-only driver
    size_t zeros_cap = 0;
    int64_t launches = 0;
    int wsplit_mask = 0xF0;      // layers whose forward-sweep weights the Jacobian kernel takes as hi + lo pairs: 4-7
                                 // (measured on 1 280 reference steps: 4-7 97.4 % within tolerance at 488 objects/s,
                                 // 1-7 96.9 % at 477, 5-7 95.9 % at 491, 1-3 96.2 % at 491, none 95.9 % at 500;
                                 // experiment knob: environment variable DSPSLAM_WSPLIT_MASK at handle creation)
    bool tc_ready = false;
    std::string tc_why;          // why the tcgen05 path is unavailable for this decoder
    std::mutex mu;
    // profiling (dspslam_profile_*): event pairs around decoder launches + device row counters
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_ev;      // pool, used in pairs
    std::vector<int> prof_kind;            // kind of each recorded pair
    std::vector<int64_t> prof_rows;        // host-known row count of the pair (stand-alone launches)
    unsigned long long* stats = nullptr;   // device: [0] GN forward rows, [1] GN Jacobian rows
    // The scratch above (ws, lat, zeros, tc_xtra1, tc_park) is shared by every call on the handle: calls are ordered
    // on the device by `done` (recorded at the end of every call, waited for by the next call when it arrives on a
    // different stream), so threads that hand in different CUDA streams cannot race on it.
    cudaEvent_t done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool has_done = false;
    // second stream of the GN driver: the Jacobian tiles that hold surface rows only run here, beside the ray pipeline
    // of the same iteration (gn_run); ev_state / ev_early order the two streams twice per iteration
    cudaStream_t aux = nullptr;
    cudaEvent_t ev_state = nullptr, ev_early = nullptr;
    int early_min_tiles = 1 << 30;   // OFF by default: measured neutral (the second stream cannot be given a lower priority
                                     // than a default-priority caller stream, so the early tiles run first instead of
                                     // filling the gaps: 489-492 against 491-499 objects/s, profiles/r02_notes.md);
                                     // environment variable DSPSLAM_JAC_EARLY=1 at handle creation switches it on (>= 16 tiles)
};

namespace {

struct Carver {   // bump allocator over the workspace (256-byte aligned pieces)
    size_t off = 0;
    size_t take(size_t bytes) {
        size_t at = off;
        off += (bytes + 255) & ~(size_t)255;
        return at;
    }
};

// Grows the scratch (never shrinks; geometric growth so a sequence of slowly growing batches re-allocates O(log)
// times).  Stream-ordered: the old block is released and the new one obtained on the calling stream, behind every
// kernel already queued on it (calls on other streams are ordered in front by StreamOrder) - no device-wide
// synchronisation on the hot path.
int ensure_ws(dspslam_handle* h, size_t bytes, cudaStream_t st) {
    if (bytes <= h->ws_cap) return 0;
    if (h->ws) {
        CK(cudaFreeAsync(h->ws, st));
        h->ws = nullptr;
        h->ws_cap = 0;
    }
    size_t want = bytes + bytes / 2;
    cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&h->ws), want, st);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        e = cudaMallocAsync(reinterpret_cast<void**>(&h->ws), bytes, st);
        want = bytes;
    }
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        char buf[128];
        snprintf(buf, sizeof buf, "workspace allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
        return fail(DSPSLAM_E_NOMEM, buf);
    }
    h->ws_cap = want;
    return 0;
}

// ---- FP32 parameter packing ----------------------------------------------------------------
struct Packer {
    std::vector<float> buf;
    size_t add(size_t n) {            // 64-float aligned, zero-filled
        size_t at = (buf.size() + 63) & ~(size_t)63;
        buf.resize(at + n, 0.f);
        return at;
    }
};

int pack_fp32(dspslam_handle* h, const float* const* W, const float* const* b) {
    DecoderDev& d = h->dec;
    const int L = d.L, in_dim = d.in_dim, h3 = d.h3, h3p = d.h3p;
    Packer P;
    size_t oWt0 = P.add((size_t)in_dim * HID);
    for (int i = 0; i < in_dim; ++i)
        for (int o = 0; o < HID; ++o) P.buf[oWt0 + (size_t)i * HID + o] = W[0][(size_t)o * in_dim + i];
    size_t oW0x4 = P.add(4 * HID);
    for (int a = 0; a < 3; ++a)
        for (int o = 0; o < HID; ++o) P.buf[oW0x4 + (size_t)a * HID + o] = W[0][(size_t)o * in_dim + L + a];
    size_t oWt[8] = {0}, oWn[8] = {0}, oB[8] = {0};
    for (int k = 1; k < 8; ++k) {
        const int K = (k == 4) ? h3p : HID;                 // padded reduction size of the forward product
        const int N = (k == 3) ? h3p : HID;                 // padded output size
        const int Kreal = (k == 4) ? h3 : HID, Nreal = (k == 3) ? h3 : HID;
        const int ld_in = HID;                              // W[k] is [Nreal][512]
        oWt[k] = P.add((size_t)K * N);
        for (int kk = 0; kk < Kreal; ++kk)
            for (int n = 0; n < Nreal; ++n) P.buf[oWt[k] + (size_t)kk * N + n] = W[k][(size_t)n * ld_in + kk];
        // natural copy for the backward sweep: rows = outputs (padded for k==3), 512 columns
        const int R = (k == 3) ? h3p : HID;
        oWn[k] = P.add((size_t)R * HID);
        for (int o = 0; o < Nreal; ++o)
            memcpy(&P.buf[oWn[k] + (size_t)o * HID], &W[k][(size_t)o * ld_in], sizeof(float) * HID);
    }
    size_t oW4in = P.add((size_t)in_dim * HID);
    for (int i = 0; i < in_dim; ++i)
        for (int n = 0; n < HID; ++n) P.buf[oW4in + (size_t)i * HID + n] = W[4][(size_t)n * HID + h3 + i];
    for (int k = 0; k < 8; ++k) {
        const int Nreal = (k == 3) ? h3 : HID;
        oB[k] = P.add(HID);
        memcpy(&P.buf[oB[k]], b[k], sizeof(float) * Nreal);
    }
    size_t ow8 = P.add(HID);
    memcpy(&P.buf[ow8], W[8], sizeof(float) * HID);
    d.b8 = b[8][0];

    CK(cudaMalloc(&h->wblob, P.buf.size() * sizeof(float)));
    CK(cudaMemcpy(h->wblob, P.buf.data(), P.buf.size() * sizeof(float), cudaMemcpyHostToDevice));
    const float* base = h->wblob;
    d.Wt0 = base + oWt0;
    d.W0x4 = base + oW0x4;
    d.W4in = base + oW4in;
    d.w8 = base + ow8;
    d.Wt[0] = nullptr;
    d.Wn[0] = nullptr;
    for (int k = 1; k < 8; ++k) { d.Wt[k] = base + oWt[k]; d.Wn[k] = base + oWn[k]; }
    for (int k = 0; k < 8; ++k) d.bias[k] = base + oB[k];
    return 0;
}

inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

int check_launch(dspslam_handle* h, int n_launch, const char* what) {
    h->launches += n_launch;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(DSPSLAM_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    return 0;
}

enum { PK_GN_FWD = 0, PK_GN_JAC = 1, PK_ROWS_FWD = 2, PK_ROWS_JAC = 3, PK_GN_REFINE = 4, PK_GN_JAC_EARLY = 5 };

// records the start event of a profiled launch group; returns the pair index or -1
int prof_begin(dspslam_handle* h, int kind, int64_t rows, cudaStream_t st) {
    if (!h->prof_on) return -1;
    const size_t pair = h->prof_kind.size();
    while (h->prof_ev.size() < 2 * (pair + 1)) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) { (void)cudaGetLastError(); return -1; }
        h->prof_ev.push_back(e);
    }
    h->prof_kind.push_back(kind);
    h->prof_rows.push_back(rows);
    cudaEventRecord(h->prof_ev[2 * pair], st);
    return (int)pair;
}
void prof_end(dspslam_handle* h, int pair, cudaStream_t st) {
    if (pair >= 0) cudaEventRecord(h->prof_ev[2 * pair + 1], st);
}

// Orders the calls of one handle on the device across streams (see dspslam_handle::done).
struct StreamOrder {
    dspslam_handle* h;
    cudaStream_t st;
    StreamOrder(dspslam_handle* h_, void* stream) : h(h_), st(reinterpret_cast<cudaStream_t>(stream)) {
        if (h->has_done && h->last_stream != st) cudaStreamWaitEvent(st, h->done, 0);
    }
    ~StreamOrder() {
        if (h->done && cudaEventRecord(h->done, st) == cudaSuccess) { h->has_done = true; h->last_stream = st; }
    }
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

extern "C" {

int dspslam_abi_version(void) { return DSPSLAM_ABI_VERSION; }
const char* dspslam_last_error(void) { return g_err.c_str(); }
int64_t dspslam_launch_count(const dspslam_handle* h) { return h ? h->launches : 0; }
int dspslam_query(const dspslam_handle* h, int what) {
    if (!h) { g_err = "null handle"; return -1; }
    switch (what) {
        case DSPSLAM_Q_WSPLIT_MASK: return h->wsplit_mask;
        case DSPSLAM_Q_JAC_EARLY: return h->early_min_tiles < (1 << 30) ? 1 : 0;
        case DSPSLAM_Q_DEVICE: return h->device;
        default: g_err = "unknown query"; return -1;
    }
}
int dspslam_mode_available(const dspslam_handle* h, int mode) {
    if (!h) { g_err = "null handle"; return 0; }
    if (mode == DSPSLAM_MODE_FP32) return 1;
    if (mode != DSPSLAM_MODE_TC && mode != DSPSLAM_MODE_TC_FAST && mode != DSPSLAM_MODE_TC_EXACT) { g_err = "unknown mode"; return 0; }
    if (!h->tc_ready) { g_err = h->tc_why; return 0; }
    return 1;
}

int dspslam_create(const dspslam_decoder_desc* desc, const float* const* W_host, const float* const* b_host,
                   int device, dspslam_handle** out) {
    if (!desc || !W_host || !b_host || !out) return fail(DSPSLAM_E_BADARG, "null argument");
    if (desc->hidden != HID || desc->n_hidden != 8 || desc->latent_in != 4 || desc->code_len < 1 ||
        desc->code_len > MAX_L)
        return fail(DSPSLAM_E_UNSUPPORTED,
                    "supported decoder family: 8 hidden layers of 512, latent_in=[4], code_len 1..64");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(DSPSLAM_E_BADARG, "bad device index");
    dspslam_handle* h = new (std::nothrow) dspslam_handle();
    if (!h) return fail(DSPSLAM_E_NOMEM, "host allocation failed");
    h->device = device;
    DeviceGuard g(device);
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) { delete h; return fail(DSPSLAM_E_CUDA, cudaGetErrorString(e)); }
    if (prop.major != 10) {
        delete h;
        return fail(DSPSLAM_E_UNSUPPORTED, "libdspslam_sm100 needs a Blackwell (sm_100) device");
    }
    h->sm_count = prop.multiProcessorCount;
    h->dec.L = desc->code_len;
    h->dec.in_dim = desc->code_len + 3;
    h->dec.h3 = HID - h->dec.in_dim;
    h->dec.h3p = (h->dec.h3 + 15) & ~15;
    int rc = pack_fp32(h, W_host, b_host);
    if (rc) { dspslam_destroy(h); return rc; }
    e = simt_configure();
    if (e != cudaSuccess) { dspslam_destroy(h); return fail(DSPSLAM_E_CUDA, std::string("simt_configure: ") + cudaGetErrorString(e)); }
    e = mc_configure();
    if (e != cudaSuccess) { dspslam_destroy(h); return fail(DSPSLAM_E_CUDA, std::string("mc_configure: ") + cudaGetErrorString(e)); }
    e = cudaMalloc(&h->lat, sizeof(float) * (2 * HID + MAX_L + 64));
    if (e != cudaSuccess) { dspslam_destroy(h); return fail(DSPSLAM_E_CUDA, cudaGetErrorString(e)); }
    e = cudaMalloc(&h->stats, sizeof(unsigned long long) * 8);
    if (e == cudaSuccess) e = cudaMemset(h->stats, 0, sizeof(unsigned long long) * 8);
    if (e != cudaSuccess) { dspslam_destroy(h); return fail(DSPSLAM_E_CUDA, cudaGetErrorString(e)); }
    e = cudaEventCreateWithFlags(&h->done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_state, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_early, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->aux, cudaStreamNonBlocking);
    if (e != cudaSuccess) { dspslam_destroy(h); return fail(DSPSLAM_E_CUDA, cudaGetErrorString(e)); }
    // experiment knobs (include/dspslam.h, "Environment variables"); read back with dspslam_query
    if (const char* m = getenv("DSPSLAM_JAC_EARLY")) h->early_min_tiles = (atoi(m) != 0) ? 16 : (1 << 30);
    rc = tc_pack_and_configure(h->dec, W_host, b_host, &h->hblob);
    if (rc > 1) { std::string m = tc_last_error(); dspslam_destroy(h); return fail(DSPSLAM_E_CUDA, "tcgen05 path setup: " + m); }
    h->tc_ready = (rc == 0);
    if (const char* m = getenv("DSPSLAM_WSPLIT_MASK")) {
        char* end = nullptr;
        const long v = strtol(m, &end, 0);
        if (end == m || *end != '\0' || v < 0 || v > 0xFF) {     // a typo must not silently select the tc_fast arithmetic
            dspslam_destroy(h);
            return fail(DSPSLAM_E_BADARG, std::string("DSPSLAM_WSPLIT_MASK is not an integer in 0..255: '") + m + "'");
        }
        h->wsplit_mask = (int)v & 0xFE;
    }
    if (rc == 1) h->tc_why = tc_last_error();
    *out = h;
    return 0;
}

void dspslam_destroy(dspslam_handle* h) {
    if (!h) return;
    {
        DeviceGuard g(h->device);
        cudaDeviceSynchronize();
        if (h->wblob) cudaFree(h->wblob);
        if (h->hblob) cudaFree(h->hblob);
        if (h->ws) cudaFree(h->ws);
        if (h->done) cudaEventDestroy(h->done);
        if (h->ev_state) cudaEventDestroy(h->ev_state);
        if (h->ev_early) cudaEventDestroy(h->ev_early);
        if (h->aux) cudaStreamDestroy(h->aux);
        if (h->lat) cudaFree(h->lat);
        if (h->zeros) cudaFree(h->zeros);
        if (h->stats) cudaFree(h->stats);
        for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
    }
    delete h;
}

static int run_sdf(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz, int64_t n,
                   float* sdf_out, float* jac_out) {
    if (!h || n < 0) return fail(DSPSLAM_E_BADARG, "null/negative argument");
    if (mode != DSPSLAM_MODE_FP32 && mode != DSPSLAM_MODE_TC && mode != DSPSLAM_MODE_TC_FAST && mode != DSPSLAM_MODE_TC_EXACT)
        return fail(DSPSLAM_E_BADARG, "bad mode");
    if (n == 0) return 0;                       // empty query: nothing to do (pointers may be null)
    if (!code || !xyz || !sdf_out) return fail(DSPSLAM_E_BADARG, "null argument");
    if (n > ((int64_t)1 << 31) - 256) return fail(DSPSLAM_E_BADARG, "n too large for one call");
    if (mode != DSPSLAM_MODE_FP32 && !h->tc_ready) return fail(DSPSLAM_E_UNSUPPORTED, "tcgen05 path unavailable: " + h->tc_why);
    float* c0 = h->lat;
    float* c4 = h->lat + HID;
    launch_latent_bias(h->dec, code, 1, h->dec.L, c0, c4, S(stream));
    const int pp = prof_begin(h, jac_out ? PK_ROWS_JAC : PK_ROWS_FWD, n, S(stream));
    int nl = 1;
    if (mode == DSPSLAM_MODE_FP32) {
        if (jac_out) launch_sdf_jac_simt(h->dec, xyz, n, c0, c4, sdf_out, jac_out, S(stream));
        else launch_sdf_rows_simt(h->dec, xyz, n, c0, c4, sdf_out, S(stream));
        ++nl;
    } else {
        if (jac_out && mode == DSPSLAM_MODE_TC_EXACT)
            nl += tc_launch_sdf_jac_exact(h->dec, h->sm_count, xyz, n, c0, c4, sdf_out, jac_out, S(stream));
        else if (jac_out) nl += tc_launch_sdf_jac(h->dec, h->sm_count, xyz, n, c0, c4, sdf_out, jac_out,
                                                  mode == DSPSLAM_MODE_TC ? h->wsplit_mask : 0, S(stream));
        else nl += tc_launch_sdf_rows(h->dec, h->sm_count, xyz, n, 0, 0, 0, c0, c4, sdf_out, S(stream));
    }
    prof_end(h, pp, S(stream));
    return check_launch(h, nl, "dspslam_sdf");
}

int dspslam_sdf(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz, int64_t n,
                float* sdf_out) {
    if (!h) return fail(DSPSLAM_E_BADARG, "null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    return run_sdf(h, stream, mode, code, xyz, n, sdf_out, nullptr);
}

int dspslam_sdf_jac(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz, int64_t n,
                    float* sdf_out, float* jac_out) {
    if (!h || !jac_out) return fail(DSPSLAM_E_BADARG, "null handle or jac_out");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    return run_sdf(h, stream, mode, code, xyz, n, sdf_out, jac_out);
}

int dspslam_mean_sdf(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz, int64_t n,
                     float* mean_out) {
    if (!h || !mean_out || n <= 0) return fail(DSPSLAM_E_BADARG, "bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    int rc = ensure_ws(h, (size_t)n * sizeof(float), S(stream));
    if (rc) return rc;
    float* tmp = reinterpret_cast<float*>(h->ws);
    rc = run_sdf(h, stream, mode, code, xyz, n, tmp, nullptr);
    if (rc) return rc;
    launch_mean(tmp, n, mean_out, S(stream));
    return check_launch(h, 1, "dspslam_mean_sdf");
}

int dspslam_grid_sdf(dspslam_handle* h, void* stream, int mode, const float* code, int G, int64_t i0,
                     int64_t i1, int coord_mode, float* out) {
    if (!h || !code || !out) return fail(DSPSLAM_E_BADARG, "null argument");
    if (G < 2 || G > 1024) return fail(DSPSLAM_E_BADARG, "G out of range [2,1024]");
    const int64_t total = (int64_t)G * G * G;
    if (i0 < 0 || i1 < i0 || i1 > total) return fail(DSPSLAM_E_BADARG, "bad index range");
    if (coord_mode != DSPSLAM_GRID_REFERENCE_SHEAR && coord_mode != DSPSLAM_GRID_LATTICE)
        return fail(DSPSLAM_E_BADARG, "bad coord_mode");
    if (mode != DSPSLAM_MODE_FP32 && mode != DSPSLAM_MODE_TC && mode != DSPSLAM_MODE_TC_FAST && mode != DSPSLAM_MODE_TC_EXACT)
        return fail(DSPSLAM_E_BADARG, "bad mode");
    if (i1 == i0) return 0;
    if (mode != DSPSLAM_MODE_FP32 && !h->tc_ready) return fail(DSPSLAM_E_UNSUPPORTED, "tcgen05 path unavailable: " + h->tc_why);
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    float* c0 = h->lat;
    float* c4 = h->lat + HID;
    launch_latent_bias(h->dec, code, 1, h->dec.L, c0, c4, S(stream));
    const int pp = prof_begin(h, PK_ROWS_FWD, i1 - i0, S(stream));
    int nl = 2;
    if (mode == DSPSLAM_MODE_FP32) launch_grid_simt(h->dec, G, i0, i1, coord_mode, c0, c4, out, S(stream));
    else nl = 1 + tc_launch_sdf_rows(h->dec, h->sm_count, nullptr, i1 - i0, i0, G, 1 + coord_mode, c0, c4, out, S(stream));
    prof_end(h, pp, S(stream));
    return check_launch(h, nl, "dspslam_grid_sdf");
}

int dspslam_marching_cubes(dspslam_handle* h, void* stream, const float* sdf, int G, float level, float spacing,
                           float origin, int max_verts, int max_tris, float* verts_out, int32_t* faces_out,
                           int32_t* counts_out) {
    if (!h || !sdf || !counts_out) return fail(DSPSLAM_E_BADARG, "null argument");
    if (G < 2 || G > 1024) return fail(DSPSLAM_E_BADARG, "G out of range [2,1024]");
    if (max_verts < 0 || max_tris < 0) return fail(DSPSLAM_E_BADARG, "negative capacity");
    if ((max_verts > 0 && !verts_out) || (max_tris > 0 && !faces_out)) return fail(DSPSLAM_E_BADARG, "null output");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    int rc = ensure_ws(h, mc_scratch_bytes(G), S(stream));
    if (rc) return rc;
    const int nl = launch_marching_cubes(sdf, G, level, spacing, origin, max_verts, max_tris, verts_out, faces_out,
                                         counts_out, h->ws, S(stream));
    return check_launch(h, nl, "dspslam_marching_cubes");
}

static int gn_run(dspslam_handle* h, void* stream, int mode, const JointParamsDev& prm, int iters, int ray_phases,
                  float refine_window, int B,
                  const int32_t* pts_off, const int32_t* ray_off, const int32_t* n_fg, const float* pts,
                  const float* rays, const float* depth, int64_t total_pts, int64_t total_rays, int max_pts,
                  int max_rays, const float* pose_in, float* pose_out, const float* code_in, float* code_out,
                  const float* scale_in, int32_t* status_out, float* loss_out, float* dbg_out, int dbg_stride) {
    // diagnostic mixes (not in the public header): 2 = tensor-core forward samples + FP32 Jacobian rows, 3 = the reverse
    if (mode < 0 || mode > 5) return fail(DSPSLAM_E_BADARG, "bad mode");
    if (mode != DSPSLAM_MODE_FP32 && !h->tc_ready) return fail(DSPSLAM_E_UNSUPPORTED, "tcgen05 path unavailable: " + h->tc_why);
    const bool fast = (mode == DSPSLAM_MODE_TC_FAST), exact = (mode == DSPSLAM_MODE_TC_EXACT);
    const bool fwd_tc = (mode == DSPSLAM_MODE_TC || mode == 2 || fast || exact);
    const bool jac_tc = (mode == DSPSLAM_MODE_TC || mode == 3 || fast || exact);
    if (B <= 0 || B > 65535) return fail(DSPSLAM_E_BADARG, "B must be in 1..65535");
    if (iters < 0) return fail(DSPSLAM_E_BADARG, "negative iteration count");
    const int D = prm.D;
    if (!prm.pose_only && (D < 2 || D > DSP_MAX_D)) return fail(DSPSLAM_E_BADARG, "num_depth_samples must be 2..64");
    const int64_t V = total_rays * D;
    if (total_pts + V > ((int64_t)1 << 31) - 1024) return fail(DSPSLAM_E_BADARG, "batch too large: split it");
    cudaStream_t st = S(stream);

    GnWork w{};
    w.dec = h->dec;
    w.prm = prm;
    w.B = B;
    w.total_rays = (int)total_rays;
    w.total_pts = (int)total_pts;
    w.fgroups_max = (int)(((int64_t)max_rays * D + GROUP_ROWS - 1) / GROUP_ROWS);
    w.jac_rows = exact ? REFINE_ROWS : GROUP_ROWS;
    w.jgroups_max = (int)(((int64_t)max_pts + (int64_t)max_rays * D + w.jac_rows - 1) / w.jac_rows);
    if (w.jgroups_max < 1) w.jgroups_max = 1;
    w.pts_off = pts_off; w.ray_off = ray_off; w.n_fg = n_fg;
    w.pts = pts; w.rays = rays; w.depth = depth;
    // Ray phases: 0 = choose.  Phases pay off once the batch fills the GPU several times over (every phase is
    // at least one wave of 128-row tiles); a small batch runs one pass over all in-ball samples.  Two phases
    // measured best on the C2 batch (B200: 1 phase 303, 2: 494, 3: 480, 4: 473 objects/s).
    if (ray_phases < 0 || ray_phases > MAX_PHASES) return fail(DSPSLAM_E_BADARG, "ray_phases must be 0..4");
    w.n_phases = ray_phases ? ray_phases : ((V / GROUP_ROWS >= (int64_t)h->sm_count * 8) ? 2 : 1);
    w.phase = 0;

    // second, FP32-grade evaluation of the ray samples next to a decision threshold (tensor-core forward pass only)
    const bool refine = fwd_tc && !fast && !prm.pose_only && refine_window >= 0.f;
    w.jac_wsplit = (jac_tc && !fast) ? h->wsplit_mask : 0;
    // default window: 0.05 (the decisions only); the exact mode takes 1.0 - every sample inside the occupancy band, whose
    // VALUE enters a residual - which is what brings its free-running end state to the reference's own reproducibility
    // floor (profiles/r03_refine_window_sweep.jsonl: 0.20 -> 0.47 of 128 objects, floor 0.51; 328 -> 250 objects/s)
    w.refine_eps = refine ? (refine_window > 0.f ? refine_window : (exact ? 1.0f : 0.05f)) * prm.th : 0.f;
    const size_t rgroups_max = refine ? (size_t)(((int64_t)max_rays * D + REFINE_ROWS - 1) / REFINE_ROWS) : 0;
    size_t max_tiles = (size_t)B * (size_t)(w.jgroups_max > w.fgroups_max ? w.jgroups_max : w.fgroups_max) + 1;
    if ((size_t)B * rgroups_max + 1 > max_tiles) max_tiles = (size_t)B * rgroups_max + 1;
    Carver c;
    const size_t o_st = c.take(sizeof(ObjState) * B);
    const size_t o_code = c.take(sizeof(float) * MAX_L * B);
    const size_t o_c0 = c.take(sizeof(float) * HID * B);
    const size_t o_c4 = c.take(sizeof(float) * HID * B);
    const size_t o_frows = c.take(sizeof(float4) * (size_t)V);
    const size_t o_sdf = c.take(sizeof(float) * (size_t)V);
    const size_t o_rrows = c.take(refine ? sizeof(float4) * (size_t)V : 0);
    const size_t o_rcnt = c.take(sizeof(int) * (size_t)total_rays);
    const size_t o_roff = c.take(sizeof(int) * (size_t)total_rays);
    const size_t o_rhit = c.take(sizeof(int) * (size_t)total_rays);
    const size_t o_rres = c.take(sizeof(float) * (size_t)total_rays);
    const size_t o_dtab = c.take(sizeof(float) * 64 * (size_t)B);
    const size_t o_dobj = c.take(sizeof(unsigned int) * (size_t)B);
    const int ray_chunks = max_rays > 0 ? (max_rays + 127) / 128 : 1;
    const size_t o_cc = c.take(sizeof(int) * (size_t)B * ray_chunks);
    const size_t o_jrows = c.take(sizeof(float4) * (size_t)(total_pts + V));
    const size_t o_jres = c.take(sizeof(float) * (size_t)(total_pts + V));
    const size_t o_act = c.take((size_t)total_pts);
    const size_t o_sres = c.take(sizeof(float) * (size_t)total_pts);
    const size_t o_tobj = c.take(sizeof(int) * max_tiles);
    const size_t o_trow = c.take(sizeof(int) * max_tiles);
    const size_t o_ntl = c.take(sizeof(int) * 8);            // n_tiles[4] | done[4]
    // early Jacobian tiles: the k_tc<5> launches of the joint driver, batches with enough surface-only tiles
    const size_t egroups_max = (size_t)max_pts / GROUP_ROWS;
    const bool early = jac_tc && !exact && !prm.pose_only && h->aux && iters > 0 &&
                       (size_t)B * egroups_max >= (size_t)h->early_min_tiles;
    const size_t o_etobj = c.take(early ? sizeof(int) * ((size_t)B * egroups_max + 1) : 0);
    const size_t o_etrow = c.take(early ? sizeof(int) * ((size_t)B * egroups_max + 1) : 0);
    const size_t o_toff = c.take(sizeof(int) * ((size_t)B + 1));
    const size_t o_part = c.take(sizeof(float) * PART_STRIDE * (size_t)B * w.jgroups_max);
    const size_t o_hred = c.take(sizeof(double) * PART_STRIDE * (size_t)B);
    const size_t o_xtra = c.take(mode != DSPSLAM_MODE_FP32 ? tc_xtra_bytes_per_object() * (size_t)B : 0);
    if (c.off > ((size_t)150 << 30)) return fail(DSPSLAM_E_NOMEM, "batch needs more than 150 GB of scratch: split it");
    int rc = ensure_ws(h, c.off, st);
    if (rc) return rc;
    char* base = h->ws;
    w.st = reinterpret_cast<ObjState*>(base + o_st);
    w.code = reinterpret_cast<float*>(base + o_code);
    w.c0 = reinterpret_cast<float*>(base + o_c0);
    w.c4 = reinterpret_cast<float*>(base + o_c4);
    w.frows = reinterpret_cast<float4*>(base + o_frows);
    w.sdf_samp = reinterpret_cast<float*>(base + o_sdf);
    w.rrows = reinterpret_cast<float4*>(base + o_rrows);
    w.ray_cnt = reinterpret_cast<int*>(base + o_rcnt);
    w.ray_roff = reinterpret_cast<int*>(base + o_roff);
    w.ray_hit = reinterpret_cast<int*>(base + o_rhit);
    w.ray_res = reinterpret_cast<float*>(base + o_rres);
    w.dtab = reinterpret_cast<float*>(base + o_dtab);
    w.done_obj = reinterpret_cast<unsigned int*>(base + o_dobj);
    w.chunk_cnt = reinterpret_cast<int*>(base + o_cc);
    w.ray_chunks = ray_chunks;
    w.jrows = reinterpret_cast<float4*>(base + o_jrows);
    w.jres = reinterpret_cast<float*>(base + o_jres);
    w.active = reinterpret_cast<uint8_t*>(base + o_act);
    w.sres = reinterpret_cast<float*>(base + o_sres);
    w.tile_obj = reinterpret_cast<int*>(base + o_tobj);
    w.tile_row0 = reinterpret_cast<int*>(base + o_trow);
    w.n_tiles = reinterpret_cast<int*>(base + o_ntl);
    w.done = reinterpret_cast<unsigned int*>(base + o_ntl) + 4;
    w.tile_off = reinterpret_cast<int*>(base + o_toff);
    w.early = early ? 1 : 0;
    w.etile_obj = reinterpret_cast<int*>(base + o_etobj);
    w.etile_row0 = reinterpret_cast<int*>(base + o_etrow);
    w.tc = (mode != DSPSLAM_MODE_FP32) ? 1 : 0;
    w.partials = reinterpret_cast<float*>(base + o_part);
    w.hred = reinterpret_cast<double*>(base + o_hred);
    w.stats = h->prof_on ? h->stats : nullptr;
    w.tc_xtra = reinterpret_cast<uint8_t*>(base + o_xtra);

    // Launch sequence per iteration (no host synchronisation anywhere): [sample -> forward decoder] x phases ->
    // rays -> SDF+Jacobian decoder -> solve.  The tile lists of the decoder kernels are planned by the last CTA of
    // the kernel in front of them; solve also prepares the next iteration (pose scalars, latent biases, rows).
    int nl = 0;
    CK(cudaMemsetAsync(base + o_ntl, 0, sizeof(int) * 8, st));
    CK(cudaMemsetAsync(base + o_dobj, 0, sizeof(unsigned int) * (size_t)B, st));
    launch_gn_init(w, pose_in, code_in, h->dec.L, scale_in, st); ++nl;
    if (iters > 0) { launch_gn_prep(w, 0, st); ++nl; }
    for (int e = 0; e < iters; ++e) {
        if (early) {        // state of iteration e is final (prep / solve are queued): the surface-only tiles start now
            CK(cudaEventRecord(h->ev_state, st));
            CK(cudaStreamWaitEvent(h->aux, h->ev_state, 0));
            const int pe = prof_begin(h, PK_GN_JAC_EARLY, 0, h->aux);
            nl += tc_launch_gn_jac(w, h->sm_count, 1, h->aux);
            prof_end(h, pe, h->aux);
            CK(cudaEventRecord(h->ev_early, h->aux));
        }
        if (!prm.pose_only) {
            for (int ph = 0; ph < w.n_phases; ++ph) {      // front-to-back ray phases (k_gn_sample)
                w.phase = ph;
                launch_gn_sample(w, st); ++nl;
                const int pf = prof_begin(h, PK_GN_FWD, 0, st);
                if (!fwd_tc) { launch_gn_fwd_simt(w, st); ++nl; }
                else nl += tc_launch_gn_fwd(w, h->sm_count, st);
                prof_end(h, pf, st);
            }
            w.phase = 0;
            if (refine) {
                launch_gn_refine_plan(w, st); ++nl;
                const int pr = prof_begin(h, PK_GN_REFINE, 0, st);
                nl += tc_launch_gn_refine(w, h->sm_count, st);
                prof_end(h, pr, st);
            }
            launch_gn_rays(w, max_rays, st); nl += 2;
        }
        const int pj = prof_begin(h, PK_GN_JAC, 0, st);
        if (!jac_tc) { launch_gn_jac_simt(w, st); ++nl; }
        else if (exact) nl += tc_launch_gn_jac_exact(w, h->sm_count, st);
        else nl += tc_launch_gn_jac(w, h->sm_count, 0, st);
        prof_end(h, pj, st);
        if (early) CK(cudaStreamWaitEvent(st, h->ev_early, 0));     // solve reads the early tiles' partials
        launch_gn_solve(w, e, e == iters - 1, dbg_out, dbg_stride, st); nl += 2;
    }
    launch_gn_final(w, pose_out, code_out, h->dec.L, scale_in, status_out, loss_out, st); ++nl;
    return check_launch(h, nl, "dspslam_gn");
}

int dspslam_gn_joint(dspslam_handle* h, void* stream, int mode, const dspslam_joint_params* p, int B,
                     const int32_t* pts_off, const int32_t* ray_off, const int32_t* n_fg, const float* pts,
                     const float* rays, const float* depth, int64_t total_pts, int64_t total_rays, int max_pts,
                     int max_rays, float* t_cam_obj_io, float* code_io, int32_t* status_out, float* loss_out,
                     float* dbg_out) {
    if (!h || !p || !pts_off || !ray_off || !n_fg || !t_cam_obj_io || !code_io)
        return fail(DSPSLAM_E_BADARG, "null argument");
    if ((total_pts > 0 && !pts) || (total_rays > 0 && (!rays || !depth))) return fail(DSPSLAM_E_BADARG, "null data array");
    if (total_pts < 0 || total_rays < 0 || max_pts < 0 || max_rays < 0) return fail(DSPSLAM_E_BADARG, "negative size");
    JointParamsDev prm{};
    prm.k1 = p->k1; prm.k2 = p->k2; prm.k3 = p->k3; prm.k4 = p->k4;
    prm.b1 = p->b1; prm.b2 = p->b2; prm.lr = p->lr; prm.s_damp = p->s_damp; prm.th = p->cut_off;
    prm.D = p->num_depth_samples;
    prm.pose_only = 0;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    return gn_run(h, stream, mode, prm, p->num_iterations, p->ray_phases, p->refine_window, B, pts_off, ray_off, n_fg, pts, rays, depth, total_pts,
                  total_rays, max_pts, max_rays, t_cam_obj_io, t_cam_obj_io, code_io, code_io, nullptr, status_out,
                  loss_out, dbg_out, DSPSLAM_DBG_STRIDE);
}

int dspslam_gn_pose_only(dspslam_handle* h, void* stream, int mode, int num_iterations, int B,
                         const int32_t* pts_off, const float* pts, int64_t total_pts, int max_pts,
                         const float* scale, const float* code, float* t_co_se3_io, float* dbg_out) {
    if (!h || !pts_off || !pts || !scale || !code || !t_co_se3_io) return fail(DSPSLAM_E_BADARG, "null argument");
    if (total_pts <= 0 || max_pts <= 0 || B <= 0) return fail(DSPSLAM_E_BADARG, "no surface points");
    JointParamsDev prm{};
    prm.pose_only = 1;
    prm.D = 0;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    // the kernels index ray_off / n_fg even without rays: hand them zero arrays
    if (h->zeros_cap < (size_t)B + 1) {
        if (h->zeros) { CK(cudaFreeAsync(h->zeros, S(stream))); h->zeros = nullptr; h->zeros_cap = 0; }
        const size_t cap = 2 * (size_t)B + 1025;
        CK(cudaMallocAsync(reinterpret_cast<void**>(&h->zeros), sizeof(int32_t) * cap, S(stream)));
        CK(cudaMemsetAsync(h->zeros, 0, sizeof(int32_t) * cap, S(stream)));
        h->zeros_cap = cap;
    }
    return gn_run(h, stream, mode, prm, num_iterations, 1, -1.f, B, pts_off, h->zeros, h->zeros, pts, nullptr, nullptr,
                  total_pts, 0, max_pts, 0, t_co_se3_io, t_co_se3_io, code, nullptr, scale, nullptr, nullptr, dbg_out,
                  64);
}

static int frame_params(const dspslam_frame_params* p, FrameParamsDev& d) {
    if (p->img_w <= 0 || p->img_h <= 0 || p->num_lidar_max <= 0 || p->downsample_ratio <= 0 || p->max_background < 0 ||
        p->max_background > 1024)
        return fail(DSPSLAM_E_BADARG, "frame parameters out of range");
    d.img_w = p->img_w; d.img_h = p->img_h; d.max_pts = p->num_lidar_max; d.min_mask_area = p->min_mask_area;
    d.alpha = p->downsample_ratio; d.max_bg = p->max_background;
    memcpy(d.K, p->K, sizeof d.K);
    memcpy(d.invK, p->invK, sizeof d.invK);
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) d.Rcv[3 * i + j] = p->T_cam_velo[4 * i + j];
        d.tcv[i] = p->T_cam_velo[4 * i + 3];
    }
    d.k1 = p->k1; d.k2 = p->k2;
    memcpy(d.Kd, p->K64, sizeof d.Kd);
    memcpy(d.invKd, p->invK64, sizeof d.invKd);
    return 0;
}

int dspslam_frame_build_kitti(dspslam_handle* h, void* stream, const dspslam_frame_params* prm, const float* velo,
                              int64_t n_velo, const dspslam_frame_box* boxes_host, int B, const uint8_t* masks,
                              int n_masks, const float* bboxes, int32_t* pts_off, int32_t* ray_off, int32_t* n_fg,
                              float* pts, float* rays, float* depth, int32_t* info) {
    if (!h || !prm || !boxes_host || !pts_off || !ray_off || !n_fg || !pts || !rays || !depth || !info)
        return fail(DSPSLAM_E_BADARG, "null argument");
    if (B <= 0 || B > FB_MAX_DETS) return fail(DSPSLAM_E_BADARG, "B out of range [1,256]");
    if (n_velo < 0 || (n_velo > 0 && !velo)) return fail(DSPSLAM_E_BADARG, "bad scan");
    if (n_masks < 0 || n_masks > FB_MAX_MASKS || (n_masks > 0 && (!masks || !bboxes)))
        return fail(DSPSLAM_E_BADARG, "n_masks out of range [0,256] or null masks");
    if ((reinterpret_cast<uintptr_t>(velo) & 15) != 0) return fail(DSPSLAM_E_BADARG, "scan must be 16-byte aligned");
    FrameParamsDev d{};
    int rc = frame_params(prm, d);
    if (rc) return rc;
    std::vector<FrameBoxDev> bx((size_t)B);
    std::vector<int> front((size_t)B);
    for (int i = 0; i < B; ++i) {
        const dspslam_frame_box& s = boxes_host[i];
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) bx[i].Rov[3 * r + c] = s.T_obj_velo[4 * r + c];
            bx[i].tov[r] = s.T_obj_velo[4 * r + 3];
            bx[i].lo[r] = s.lo[r]; bx[i].hi[r] = s.hi[r]; bx[i].half[r] = s.half[r];
        }
        front[i] = s.is_front;
    }
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    rc = ensure_ws(h, frame_scratch_bytes(n_velo, B, n_masks, d.max_bg), S(stream));
    if (rc) return rc;
    const int nl = launch_frame_kitti(d, velo, n_velo, bx.data(), front.data(), B, masks, n_masks, bboxes, pts_off,
                                      ray_off, n_fg, pts, rays, depth, info, h->ws, S(stream));
    return check_launch(h, nl, "dspslam_frame_build_kitti");
}

int dspslam_frame_build_mono(dspslam_handle* h, void* stream, const dspslam_frame_params* prm, const uint8_t* masks,
                             int n_masks, const float* bboxes, float* rays_out, int32_t* info_out) {
    if (!h || !prm || !masks || !bboxes || !rays_out || !info_out) return fail(DSPSLAM_E_BADARG, "null argument");
    if (n_masks <= 0 || n_masks > FB_MAX_MASKS) return fail(DSPSLAM_E_BADARG, "n_masks out of range [1,256]");
    FrameParamsDev d{};
    int rc = frame_params(prm, d);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    rc = ensure_ws(h, frame_scratch_bytes(0, 1, n_masks, d.max_bg), S(stream));
    if (rc) return rc;
    const int nl = launch_frame_mono(d, masks, n_masks, bboxes, rays_out, info_out, h->ws, S(stream));
    return check_launch(h, nl, "dspslam_frame_build_mono");
}

int dspslam_pg_from_matrices(dspslam_handle* h, void* stream, const float* T, int n, double* poses_out) {
    if (!h || (n > 0 && (!T || !poses_out))) return fail(DSPSLAM_E_BADARG, "null argument");
    if (n < 0) return fail(DSPSLAM_E_BADARG, "negative count");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    return check_launch(h, launch_pg_from_matrices(T, n, poses_out, S(stream)), "dspslam_pg_from_matrices");
}

int dspslam_pg_linearize(dspslam_handle* h, void* stream, const double* vertices, int n_vertices,
                         const int32_t* edge_i, const int32_t* edge_j, const double* measurements, int n_edges,
                         double inv_sigma, double huber_delta, double* error, double* Ji, double* Jj, double* chi2,
                         double* rho, double* weight, double* A, double* b, double* Hij) {
    if (!h) return fail(DSPSLAM_E_BADARG, "null handle");
    if (n_vertices < 0 || n_edges < 0) return fail(DSPSLAM_E_BADARG, "negative count");
    if (n_edges > 0 && (!vertices || !edge_i || !edge_j || !measurements || !error || !Ji || !Jj || !chi2 || !rho ||
                        !weight))
        return fail(DSPSLAM_E_BADARG, "null argument");
    if ((A == nullptr) != (b == nullptr)) return fail(DSPSLAM_E_BADARG, "A and b go together");
    if (!(inv_sigma > 0.0)) return fail(DSPSLAM_E_BADARG, "inv_sigma must be positive");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    const int nl = launch_pg_linearize(vertices, n_vertices, edge_i, edge_j, measurements, n_edges, inv_sigma,
                                       huber_delta, error, Ji, Jj, chi2, rho, weight, A, b, Hij, S(stream));
    return check_launch(h, nl, "dspslam_pg_linearize");
}

int dspslam_pg_oplus(dspslam_handle* h, void* stream, double* vertices, const double* updates, int n) {
    if (!h || (n > 0 && (!vertices || !updates))) return fail(DSPSLAM_E_BADARG, "null argument");
    if (n < 0) return fail(DSPSLAM_E_BADARG, "negative count");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    StreamOrder so(h, stream);
    return check_launch(h, launch_pg_oplus(vertices, updates, n, S(stream)), "dspslam_pg_oplus");
}

#ifdef DSP_TC_TRACE
// debug builds only (tools/trace_jac.py): the event log CTA 0 of the last GN Jacobian launch left in the park buffer
int dspslam_debug_trace(dspslam_handle* h, long long* out_host, int n) {
    DeviceGuard g(h->device);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out_host, h->dec.tc_park + (size_t)200 * 128 * 68, sizeof(long long) * (size_t)n, cudaMemcpyDeviceToHost));
    return 0;
}
#endif

int dspslam_profile_enable(dspslam_handle* h, int on) {
    if (!h) return fail(DSPSLAM_E_BADARG, "null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    CK(cudaDeviceSynchronize());
    h->prof_on = on != 0;
    h->prof_kind.clear();
    h->prof_rows.clear();
    CK(cudaMemset(h->stats, 0, sizeof(unsigned long long) * 8));
    return 0;
}

int dspslam_profile_read(dspslam_handle* h, dspslam_profile* out) {
    if (!h || !out) return fail(DSPSLAM_E_BADARG, "null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    memset(out, 0, sizeof(*out));
    double ms[6] = {0, 0, 0, 0, 0, 0};
    int64_t rows[6] = {0, 0, 0, 0, 0, 0}, cnt[6] = {0, 0, 0, 0, 0, 0};
    for (size_t p = 0; p < h->prof_kind.size(); ++p) {
        CK(cudaEventSynchronize(h->prof_ev[2 * p + 1]));
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, h->prof_ev[2 * p], h->prof_ev[2 * p + 1]));
        const int k = h->prof_kind[p];
        ms[k] += t;
        rows[k] += h->prof_rows[p];
        cnt[k] += 1;
    }
    unsigned long long st[8];
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(st, h->stats, sizeof(st), cudaMemcpyDeviceToHost));
    CK(cudaMemset(h->stats, 0, sizeof(st)));
    h->prof_kind.clear();
    h->prof_rows.clear();
    // the early Jacobian launch shares the SMs with the ray pipeline on the other stream: its event time is the
    // span it was resident for, an upper bound of its own SM time
    out->ms_gn_fwd = ms[PK_GN_FWD]; out->ms_gn_jac = ms[PK_GN_JAC] + ms[PK_GN_JAC_EARLY];
    out->ms_gn_jac_early = ms[PK_GN_JAC_EARLY];
    out->ms_rows_fwd = ms[PK_ROWS_FWD]; out->ms_rows_jac = ms[PK_ROWS_JAC];
    out->rows_gn_fwd = (int64_t)st[0]; out->rows_gn_jac = (int64_t)st[1]; out->rows_gn_inball = (int64_t)st[2];
    out->rows_fwd = rows[PK_ROWS_FWD]; out->rows_jac = rows[PK_ROWS_JAC];
    out->n_gn_fwd = cnt[PK_GN_FWD]; out->n_gn_jac = cnt[PK_GN_JAC];
    out->n_rows_fwd = cnt[PK_ROWS_FWD]; out->n_rows_jac = cnt[PK_ROWS_JAC];
    out->rows_gn_refine = (int64_t)st[3]; out->n_gn_refine = cnt[PK_GN_REFINE]; out->ms_gn_refine = ms[PK_GN_REFINE];
    return 0;
}

}  // extern "C"

// ---- dspslam_mean_sdf_noalloc: dspslam_mean_sdf for callers that must not allocate device memory ------------------
// dspslam_mean_sdf parks the n decoder outputs in the handle's workspace in front of the reduction, and grows that
// workspace with cudaMallocAsync when n exceeds what earlier calls needed.  This variant parks them in a fixed array
// that is part of the library image instead, so the call itself never asks the driver for memory.
namespace {

// One copy per GPU that loads the library (1 MiB each, held until the process exits).  Shared by every handle and
// every stream of that GPU; it keeps whatever the previous call left behind.  Nothing here has to start at zero: rows
// [0, n) are overwritten by the decoder kernel before k_mean reads exactly those rows, so no clearing pass is needed.
static __device__ __align__(256) float g_noalloc_rows[DSPSLAM_NOALLOC_MAX_ROWS];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_addr[NOALLOC_MAX_DEVICES] = {nullptr};   // device address of g_noalloc_rows on each GPU, looked up once

// `g_noalloc_rows` in host code names a host-side shadow of the symbol, not the device array: handing it to a kernel
// would fault (or, on a system where the GPU can reach host memory, quietly read and write the wrong place).  The
// address a kernel can use is asked of the runtime, once per GPU, with that GPU current.
int noalloc_rows_on(int device, float** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(DSPSLAM_E_UNSUPPORTED, "device ordinal beyond the no-alloc table");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_addr[device]) {
        int cur = -1;
        CK(cudaGetDevice(&cur));
        if (cur != device) return fail(DSPSLAM_E_CUDA, "no-alloc set-up called with another device current");
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_rows));
        if (!p) return fail(DSPSLAM_E_CUDA, "cudaGetSymbolAddress returned a null address");
        g_noalloc_addr[device] = static_cast<float*>(p);
    }
    *out = g_noalloc_addr[device];
    return 0;
}

}  // namespace

extern "C" int dspslam_mean_sdf_noalloc(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz,
                                        int64_t n, float* mean_out) {
    if (!h || !mean_out || !code || !xyz || n <= 0) return fail(DSPSLAM_E_BADARG, "bad argument");
    if (n > DSPSLAM_NOALLOC_MAX_ROWS) {
        char buf[160];
        snprintf(buf, sizeof buf, "dspslam_mean_sdf_noalloc: n = %lld exceeds DSPSLAM_NOALLOC_MAX_ROWS = %d "
                 "(use dspslam_mean_sdf, which may allocate)", (long long)n, DSPSLAM_NOALLOC_MAX_ROWS);
        return fail(DSPSLAM_E_BADARG, buf);
    }
    if (mode != DSPSLAM_MODE_FP32 && mode != DSPSLAM_MODE_TC && mode != DSPSLAM_MODE_TC_FAST && mode != DSPSLAM_MODE_TC_EXACT)
        return fail(DSPSLAM_E_BADARG, "bad mode");
    if (mode != DSPSLAM_MODE_FP32 && !h->tc_ready) return fail(DSPSLAM_E_UNSUPPORTED, "tcgen05 path unavailable: " + h->tc_why);
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    float* rows = nullptr;
    int rc = noalloc_rows_on(h->device, &rows);   // first call on this GPU: symbol look-up (before any stream capture)
    if (rc) return rc;
    StreamOrder so(h, stream);
    // same launches as dspslam_mean_sdf, minus the profiling events (their pool grows on demand)
    float* c0 = h->lat;
    float* c4 = h->lat + HID;
    launch_latent_bias(h->dec, code, 1, h->dec.L, c0, c4, S(stream));
    int nl = 2;
    if (mode == DSPSLAM_MODE_FP32) launch_sdf_rows_simt(h->dec, xyz, n, c0, c4, rows, S(stream));
    else nl = 1 + tc_launch_sdf_rows(h->dec, h->sm_count, xyz, n, 0, 0, 0, c0, c4, rows, S(stream));
    launch_mean(rows, n, mean_out, S(stream));
    return check_launch(h, nl + 1, "dspslam_mean_sdf_noalloc");
}

/* libdspslam_sm100.so - C ABI of the B200-native DSP-SLAM object-reconstruction hot path.
 *
 * Every entry point replaces a piece of the reference's Python/PyTorch path
 * (paths relative to the reference repository):
 *
 *   dspslam_create / dspslam_destroy   deep_sdf/workspace.py:202-223 (decoder upload, once per process)
 *   dspslam_sdf                        reconstruct/loss_utils.py:51-79   decode_sdf
 *   dspslam_sdf_jac                    reconstruct/loss_utils.py:82-113  get_batch_sdf_jacobian
 *   dspslam_mean_sdf                   reconstruct/optimizer.py:207-213 + loss.py:46-56
 *   dspslam_grid_sdf                   reconstruct/optimizer.py:224-227 + utils.py:97-116 (voxel grid + decode_sdf)
 *   dspslam_marching_cubes             reconstruct/utils.py:119-140      convert_sdf_voxels_to_mesh (skimage marching cubes)
 *   dspslam_gn_joint                   reconstruct/optimizer.py:90-205   Optimizer.reconstruct_object
 *                                      (+ loss.py:22-43, 60-166, 169-192; loss_utils.py:176-275)
 *   dspslam_gn_pose_only               reconstruct/optimizer.py:46-87    Optimizer.estimate_pose_cam_obj
 *   dspslam_profile_enable / _read     reconstruct/loss_utils.py:278-283 get_time (the reference's only timing hook)
 *   dspslam_frame_build_kitti          reconstruct/kitti_sequence.py:136-155,180-209 (+ :70-97, loss_utils.py:23-37)
 *                                      FrameWithLiDAR.get_detections: LiDAR crop, sub-sampling, mask association,
 *                                      background pixels, rays - output in the batch layout of dspslam_gn_joint
 *   dspslam_frame_build_mono           reconstruct/mono_sequence.py:95-106  background rays of the largest mask
 *   dspslam_pg_from_matrices           src/Converter.cc:37-54              Converter::toSE3Quat
 *   dspslam_pg_linearize               include/ObjectPoseGraph.h:56-88     EdgeSE3LieAlgebra::computeError / linearizeOplus,
 *                                      chi2 + Huber + quadratic form as g2o applies them to these edges
 *                                      (src/Optimizer_util.cc:199-225, 555-583, 647-656, 701-711)
 *   dspslam_pg_oplus                   g2o VertexSE3Expmap::oplusImpl (Thirdparty/g2o/g2o/types/types_six_dof_expmap.h:73-76)
 *
 * Conventions
 *   - plain C types only; all array arguments are DEVICE pointers unless the name
 *     ends in _host; `stream` is a cudaStream_t passed as void* (0 = default stream).
 *   - matrices are row-major float32; poses are 4x4 (16 floats).
 *   - every function returns 0 on success or a DSPSLAM_E_* code; the message of the
 *     last error on the calling thread is returned by dspslam_last_error().
 *   - no exception crosses the ABI; nothing is retained from caller memory after return
 *     (asynchronous work is ordered on `stream`).
 *   - a handle may be used from several host threads and with different streams; calls on one
 *     handle serialise on an internal mutex on the host (the reference serialises on the Python GIL)
 *     and are ordered on the device by an event when consecutive calls arrive on different streams
 *     (the handle's scratch is shared).
 */
#ifndef DSPSLAM_H_
#define DSPSLAM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSPSLAM_ABI_VERSION 2

enum {
    DSPSLAM_OK = 0,
    DSPSLAM_E_BADARG = 1,
    DSPSLAM_E_CUDA = 2,
    DSPSLAM_E_NOMEM = 3,
    DSPSLAM_E_UNSUPPORTED = 4
};

/* arithmetic mode of the decoder contractions */
enum {
    DSPSLAM_MODE_FP32 = 0, /* FP32 FFMA validation mode (1e-5 parity)                    */
    DSPSLAM_MODE_TC = 1,   /* tcgen05 tensor cores, FP16 operands / FP32 TMEM accumulate, with the two precision
                            * measures that keep a GN step on the FP32 reference's.  (1) In the GN drivers, ray samples
                            * next to a decision threshold are evaluated a second time at FP32 grade
                            * (joint_params.refine_window).  (2) The SDF+Jacobian kernel (dspslam_sdf_jac, GN drivers)
                            * multiplies by hi(W) + lo(W) - two MMA passes per K chunk - in the FORWARD sweep of layers
                            * 4-7 only (DSPSLAM_Q_WSPLIT_MASK, default 0xF0); layers 0-3, layer 8 and the whole
                            * backward (input-gradient) sweep are single FP16 passes.  The forward-only entry points
                            * (dspslam_sdf, dspslam_mean_sdf[_noalloc], dspslam_grid_sdf) are single-pass FP16 in EVERY
                            * tensor-core mode, this one, _FAST and _EXACT alike (SDF within 1e-3 of FP32).           */
    DSPSLAM_MODE_TC_FAST = 4, /* tcgen05, single FP16 pass everywhere (round-1 arithmetic): ~12 % faster, per-step
                               * parity with the reference 88 % instead of >= 95 % of the steps          */
    DSPSLAM_MODE_TC_EXACT = 5 /* tcgen05 at FP32 grade where it matters: the SDF+Jacobian rows (dspslam_sdf_jac, GN
                               * drivers) are evaluated with both operands as hi+lo FP16 pairs in both sweeps
                               * (k_tc_jacx, 64-row tiles; ~4x the tensor work per row), and every ray sample inside
                               * the occupancy band is evaluated a second time at FP32 grade (refine_window default 1.0
                               * instead of 0.05).  Forward-only entry points: single-pass FP16, see above.              */
};

/* per-object status written by the Gauss-Newton drivers */
enum {
    DSPSLAM_ST_OK = 0,
    DSPSLAM_ST_FEW_POINTS = 1, /* < 10 ray samples inside the unit ball (loss.py:87-88)       */
    DSPSLAM_ST_NAN = 2         /* NaN loss, e.g. no render row survived (optimizer.py:137,151) */
};

enum { DSPSLAM_GRID_REFERENCE_SHEAR = 0, DSPSLAM_GRID_LATTICE = 1 };

typedef struct dspslam_handle dspslam_handle;

/* Decoder geometry.  Supported family: in = [code(code_len) | xyz(3)], `n_hidden` = 8 layers of
 * `hidden` = 512 units, the input re-injected at layer `latent_in` = 4 (so layer 3 emits
 * hidden - (code_len+3) features), ReLU, one output with a final tanh.  */
typedef struct {
    int32_t code_len;  /* 1..64 */
    int32_t hidden;    /* 512   */
    int32_t n_hidden;  /* 8     */
    int32_t latent_in; /* 4     */
} dspslam_decoder_desc;

/* Effective (weight-norm folded) FP32 parameters on the HOST, natural nn.Linear layout
 * W[k] is [out_k][in_k] row-major, b[k] is [out_k], k = 0..n_hidden (9 layers). */
int dspslam_create(const dspslam_decoder_desc* desc, const float* const* W_host,
                   const float* const* b_host, int device, dspslam_handle** out);
void dspslam_destroy(dspslam_handle* h);
const char* dspslam_last_error(void);
int dspslam_abi_version(void);
/* 1 if `mode` can run on this handle's decoder, else 0 with the reason in dspslam_last_error(): the tcgen05 path
 * needs (512 - code_len - 3), rounded up to 16, to be a multiple of its 32-wide K stages (code_len 64 and 32 - the
 * two lengths the reference's callers handle, src/LocalMapping_util.cc:416-426 - qualify). */
int dspslam_mode_available(const dspslam_handle* h, int mode);
/* Settings a handle runs with, as integers (-1 and a message for a null handle / unknown `what`).
 * Environment variables (read ONCE, by dspslam_create; experiment knobs, not needed in production - both change what
 * DSPSLAM_MODE_TC does for every handle created afterwards in the process, so a deployment should leave them unset
 * and a benchmark should print dspslam_query's answers next to its numbers):
 *   DSPSLAM_WSPLIT_MASK  integer 0..255 (any strtol base); bit k (k = 1..7) = the SDF+Jacobian kernel of
 *                        DSPSLAM_MODE_TC multiplies by hi(W_k) + lo(W_k) in the FORWARD sweep of layer k.  Default 0xF0
 *                        (layers 4-7).  0 gives DSPSLAM_MODE_TC the Jacobian arithmetic of DSPSLAM_MODE_TC_FAST.  Anything
 *                        that does not parse makes dspslam_create fail with DSPSLAM_E_BADARG.
 *   DSPSLAM_JAC_EARLY    non-zero: Jacobian tiles that hold surface rows only run on a second stream beside the ray
 *                        pipeline (bit-identical results, measured neutral; default off). */
enum { DSPSLAM_Q_WSPLIT_MASK = 0, DSPSLAM_Q_JAC_EARLY = 1, DSPSLAM_Q_DEVICE = 2 };
int dspslam_query(const dspslam_handle* h, int what);
/* number of kernels launched by this handle so far (for bench.py's gpu_launches) */
int64_t dspslam_launch_count(const dspslam_handle* h);

/* sdf_out[i] = decoder([code, xyz[i]]),  i < n   (one code shared by all rows, as decode_sdf does) */
int dspslam_sdf(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz,
                int64_t n, float* sdf_out);

/* additionally jac_out[i][0..code_len+3) = d sdf / d [code, xyz]  (row stride code_len+3) */
int dspslam_sdf_jac(dspslam_handle* h, void* stream, int mode, const float* code, const float* xyz,
                    int64_t n, float* sdf_out, float* jac_out);

/* mean_out[0] = mean_i sdf(code, xyz[i])   (device scalar) */
int dspslam_mean_sdf(dspslam_handle* h, void* stream, int mode, const float* code,
                     const float* xyz, int64_t n, float* mean_out);

/* dspslam_mean_sdf for callers that are not allowed to allocate device memory.  No counterpart in the reference (its
 * PyTorch path allocates freely); it exists for hosts of this library that cannot: a tracking or mapping thread that
 * runs under a fixed memory budget handed out at start-up, a process whose GPU memory is owned by another allocator
 * (a memory pool or an arena the library must not bypass), or a real-time loop in which a driver allocation - a call
 * that can block and synchronise the device - is not acceptable.  dspslam_mean_sdf parks the n decoder outputs in a
 * workspace that it grows with cudaMallocAsync on demand; this variant parks them in a fixed array inside the
 * library image and makes no cudaMalloc / cudaFree / allocating library call at all.
 *   - limit: n <= DSPSLAM_NOALLOC_MAX_ROWS; a larger n is refused with DSPSLAM_E_BADARG and a message naming the limit.
 *     The array costs DSPSLAM_NOALLOC_MAX_ROWS * 4 bytes = 1 MiB on every GPU the library is loaded on, until exit.
 *   - the first call on each GPU looks the array's device address up (cudaGetSymbolAddress, with that GPU current);
 *     make that first call during set-up, before any stream capture begins.
 *   - the array is shared by all handles and streams of a GPU and keeps the previous call's contents (nothing needs
 *     clearing: the rows a call reads are the rows it has just written).  Calls on ONE handle are ordered as for every
 *     other entry point; calls on DIFFERENT handles of the same GPU that overlap on the device race on the array -
 *     put them on one stream or order them with events.
 * Same arithmetic, same result as dspslam_mean_sdf (bit-identical on the same mode). */
#define DSPSLAM_NOALLOC_MAX_ROWS 262144
int dspslam_mean_sdf_noalloc(dspslam_handle* h, void* stream, int mode, const float* code,
                             const float* xyz, int64_t n, float* mean_out);

/* SDF over linear grid indices [i0, i1) of a G^3 grid over [-1,1]^3 (x slowest, z fastest);
 * out[i - i0].  coord_mode: DSPSLAM_GRID_REFERENCE_SHEAR reproduces the reference's
 * true-division grid, DSPSLAM_GRID_LATTICE the regular lattice. */
int dspslam_grid_sdf(dspslam_handle* h, void* stream, int mode, const float* code, int G,
                     int64_t i0, int64_t i1, int coord_mode, float* out);

/* Iso-surface of a dense (G,G,G) float32 grid (x slowest, z fastest) at `level` as an indexed triangle mesh:
 * vertex = origin + spacing * (grid coordinate of the edge crossing, linear interpolation), one vertex per crossed
 * grid edge (welded), triangles wound so that their normals point towards larger values.
 * counts_out (device int32[3]) = {n_vertices, n_triangles, fits}.  The buffers are only written when
 * n_vertices <= max_verts and n_triangles <= max_tris (fits = 1); otherwise call again with larger buffers
 * (max_verts = max_tris = 0 just counts).  Deterministic order: vertices by owning node (x, y, z), then axis;
 * triangles by cell.  Replaces the CPU marching cubes of reconstruct/utils.py:127-129. */
int dspslam_marching_cubes(dspslam_handle* h, void* stream, const float* sdf, int G, float level,
                           float spacing, float origin, int max_verts, int max_tris, float* verts_out,
                           int32_t* faces_out, int32_t* counts_out);

/* Hyper-parameters of the joint optimisation (configs/config_*.json "optimizer") */
typedef struct {
    float k1, k2, k3, k4; /* render, sdf, code-prior, rotation-prior weights */
    float b1, b2;         /* Huber thresholds: render, sdf                    */
    float lr;             /* learning_rate                                    */
    float s_damp;         /* scale_damping                                    */
    float cut_off;        /* cut_off_threshold (occupancy band)               */
    int32_t num_iterations;
    int32_t num_depth_samples; /* <= 64 */
    int32_t ray_phases;        /* render term: front-to-back passes over the ray samples per iteration.
                                * 1 evaluates every in-ball sample like the reference; 2..4 stop a ray at its first
                                * sample with sdf <= -cut_off (occupancy exactly 1, transmittance exactly 0 behind
                                * it: bit-identical results, fewer decoder queries); 0 lets the library choose. */
    float refine_window;       /* DSPSLAM_MODE_TC only.  The render term takes decisions at |sdf| = cut_off (occupancy
                                * clamp and with_grad mask, loss.py:98-102).  A ray sample the FP16-operand forward pass
                                * leaves within refine_window * cut_off of such a threshold is evaluated a second time
                                * with hi+lo FP16 operand pairs (FP32-grade) before the per-ray scan, so the selected row
                                * set follows the FP32 reference.  0 = library default: 0.05 in DSPSLAM_MODE_TC (the decisions
                                * only), 1.0 in DSPSLAM_MODE_TC_EXACT (every sample inside the occupancy band, so that the
                                * VALUES entering the residuals are FP32-grade too); < 0 switches it off. */
} dspslam_joint_params;

/* A batch of B independent detections, concatenated.
 *   pts_off[B+1]  offsets (in points) into pts;   ray_off[B+1] offsets (in rays) into rays/depth
 *   n_fg[B]       number of foreground rays of each object (the first n_fg rays carry a depth)
 *   pts[sum M][3] surface points, camera frame;  rays[sum N][3];  depth[sum N] (background entries ignored)
 *   t_cam_obj_io[B][16]  in: initial object->camera Sim(3); out: optimised
 *   code_io[B][code_len] in: initial code; out: optimised
 *   status_out[B] int32, loss_out[B] float (k1*render + k2*sdf of the last iteration)
 *   dbg_out: NULL or [num_iterations][B][DSPSLAM_DBG_STRIDE] floats: H (71x71 row-major), b (71),
 *            then K, n_valid, sdf_loss, render_loss, scale, status per iteration.
 */
#define DSPSLAM_POSE_DIM 7
#define DSPSLAM_DBG_STRIDE 5120

int dspslam_gn_joint(dspslam_handle* h, void* stream, int mode, const dspslam_joint_params* prm,
                     int B, const int32_t* pts_off, const int32_t* ray_off, const int32_t* n_fg,
                     const float* pts, const float* rays, const float* depth,
                     int64_t total_pts, int64_t total_rays, int max_pts, int max_rays,
                     float* t_cam_obj_io, float* code_io, int32_t* status_out, float* loss_out,
                     float* dbg_out);

/* Pose-only optimisation: t_co_se3_io[B][16] SE(3) in/out, scale[B], code[B][code_len] fixed.
 * dbg_out: NULL or [num_iterations][B][64]: H (6x6), b (6), M_active. */
int dspslam_gn_pose_only(dspslam_handle* h, void* stream, int mode, int num_iterations, int B,
                         const int32_t* pts_off, const float* pts, int64_t total_pts, int max_pts,
                         const float* scale, const float* code, float* t_co_se3_io,
                         float* dbg_out);

/* Per-kernel-class device timing (CUDA events recorded on the caller's stream around the decoder
 * kernels) and algorithmic row counters, for bench.py's roofline.  Replaces the reference's wall-clock
 * `get_time()` prints.  Off by default; enabling it costs two event records per decoder launch.
 *   ms_*     summed event time of the forward-only decoder launches of the GN loop (ray samples), the
 *            SDF+Jacobian launches of the GN loop, and the stand-alone sdf / sdf_jac / grid launches
 *   rows_*   rows those launches were asked to evaluate (GN: ray samples evaluated; active surface + render rows)
 *   n_*      number of launches in each class
 * dspslam_profile_read synchronises the recorded events, returns the totals since the last read and resets. */
typedef struct {
    double ms_gn_fwd, ms_gn_jac, ms_rows_fwd, ms_rows_jac;
    int64_t rows_gn_fwd, rows_gn_jac, rows_fwd, rows_jac;
    int64_t n_gn_fwd, n_gn_jac, n_rows_fwd, n_rows_jac;
    int64_t rows_gn_inball; /* in-ball ray samples of the GN loop (what the reference evaluates; >= rows_gn_fwd) */
    int64_t rows_gn_refine; /* ray samples evaluated a second time at FP32 grade (joint_params.refine_window)     */
    int64_t n_gn_refine;
    double ms_gn_refine;
    double ms_gn_jac_early; /* part of ms_gn_jac: the Jacobian tiles of surface rows only, run on the handle's second
                             * stream beside the ray pipeline (event span while sharing the SMs, an upper bound) */
} dspslam_profile;
int dspslam_profile_enable(dspslam_handle* h, int on);
int dspslam_profile_read(dspslam_handle* h, dspslam_profile* out);

/* ---- detection -> optimiser-input builder (the step in front of the GN loop) ----------------------------------
 * Bit-identical to the reference's numpy (FP32 / FP64 evaluation order reproduced, stable compaction). */
typedef struct {
    int32_t img_w, img_h;
    int32_t num_lidar_max;    /* configs.num_lidar_max: surface points kept per detection (kitti_sequence.py:149-151) */
    int32_t min_mask_area;    /* configs.min_mask_area (kitti_sequence.py:196)                                       */
    int32_t downsample_ratio; /* int(configs.downsample_ratio): lattice stride of pixels_sampler (:71)              */
    int32_t max_background;   /* 200 background pixels per detection (:199-201), <= 1024                            */
    float K[9], invK[9];      /* camera intrinsics and their inverse, row-major (sequence.K_cam / invK_cam)         */
    float T_cam_velo[12];     /* first three rows of the velodyne->camera transform, row-major                      */
    double k1, k2;            /* radial distortion of the monocular camera (mono_sequence.py:103)                   */
    double K64[9], invK64[9]; /* monocular path: MonoSequence keeps FP64 intrinsics (mono_sequence.py:130-131)      */
} dspslam_frame_params;

/* Per-detection 4x4 bookkeeping done by the host mirror exactly as the reference does it in numpy
 * (kitti_sequence.py:125-135,141-143,163): inverse box pose, 3 m cube, half extents (w,l enlarged by 1.1). */
typedef struct {
    float T_obj_velo[12]; /* first three rows, row-major */
    float lo[3], hi[3];   /* trans -/+ 3 m                */
    float half[3];        /* (1.1 w/2, h/2, 1.1 l/2)     */
    int32_t is_front;     /* T_cam_obj[2,3] > 0          */
} dspslam_frame_box;

#define DSPSLAM_FRAME_INFO 8 /* int32 per detection: n_in_box, M, n_in_fov, mask_id (-1 none), has_rays, n_bg, hits, 0 */
#define DSPSLAM_FRAME_MAX_DETS 256
#define DSPSLAM_FRAME_MAX_MASKS 256

/* One KITTI stereo+LiDAR frame: B detections (depth order, boxes_host) against the scan velo[n_velo][4] and the
 * 2-D instance masks masks[n_masks][img_h][img_w] (bytes, non-zero = on) / bboxes[n_masks][4] (l,t,r,b).
 * Outputs (device, capacities: pts 3*B*num_lidar_max floats, rays 3*B*(num_lidar_max+max_background),
 * depth B*(num_lidar_max+max_background)) are the pts_off / ray_off / n_fg / pts / rays / depth arguments of
 * dspslam_gn_joint; detections without rays (not in front, no mask, mask too small) keep their surface points
 * and get zero rays.  info[B][DSPSLAM_FRAME_INFO]. */
int dspslam_frame_build_kitti(dspslam_handle* h, void* stream, const dspslam_frame_params* prm,
                              const float* velo, int64_t n_velo, const dspslam_frame_box* boxes_host, int B,
                              const uint8_t* masks, int n_masks, const float* bboxes,
                              int32_t* pts_off, int32_t* ray_off, int32_t* n_fg, float* pts, float* rays,
                              float* depth, int32_t* info);

/* Monocular frame: background rays of the largest mask (rays_out[max_background][3]); info_out[2] = {mask_id, n_bg}. */
int dspslam_frame_build_mono(dspslam_handle* h, void* stream, const dspslam_frame_params* prm,
                             const uint8_t* masks, int n_masks, const float* bboxes, float* rays_out,
                             int32_t* info_out);

/* ---- camera-object pose-graph edges (the consumer of the SE(3) measurement of dspslam_gn_pose_only) ----------
 * All arrays FP64 on the device.  A pose is g2o's Vector7d: (tx, ty, tz, qx, qy, qz, qw), w >= 0, unit norm. */

/* T[n][16] row-major FP32 4x4 -> poses_out[n][7] (Converter::toSE3Quat: double R, t -> SE3Quat(R, t)). */
int dspslam_pg_from_matrices(dspslam_handle* h, void* stream, const float* T, int n, double* poses_out);

/* For every edge e between vertices[edge_i[e]] (keyframe, T_cw) and vertices[edge_j[e]] (object, T_ow) with
 * measurement measurements[e] (T_co):
 *   error[e][6]  = log(Tij^-1 Ti Tj^-1)  (omega, upsilon)          Ji[e][6][6], Jj[e][6][6] row-major
 *   chi2[e]      = error^T (inv_sigma I) error                     rho[e], weight[e] = Huber rho, rho'
 *                  (huber_delta <= 0: no robust kernel, rho = chi2, weight = 1)
 * and, when A and b are given (both or neither; Hij optional), g2o's quadratic form with the robust information
 * weight * inv_sigma * I:  A[v][6][6] = sum J^T W J,  b[v][6] = sum J^T (-W error) over the edges of vertex v in edge
 * order,  Hij[e][6][6] = Ji^T W Jj.  n_edges = 0 is allowed (A, b are zero-filled). */
int dspslam_pg_linearize(dspslam_handle* h, void* stream, const double* vertices, int n_vertices,
                         const int32_t* edge_i, const int32_t* edge_j, const double* measurements, int n_edges,
                         double inv_sigma, double huber_delta, double* error, double* Ji, double* Jj, double* chi2,
                         double* rho, double* weight, double* A, double* b, double* Hij);

/* vertices[v] <- exp(updates[v]) * vertices[v], updates[n][6] = (omega, upsilon). */
int dspslam_pg_oplus(dspslam_handle* h, void* stream, double* vertices, const double* updates, int n);

#ifdef __cplusplus
}
#endif
#endif /* DSPSLAM_H_ */

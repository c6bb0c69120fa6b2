// Kernel-launcher interface between api.cu and the kernel translation units.
#pragma once
#include "common.cuh"

namespace dsp {

// Device view of one Gauss-Newton batch: inputs, state and scratch (all device pointers).
struct GnWork {
    DecoderDev dec;
    JointParamsDev prm;
    int B;
    int total_rays, total_pts;
    int jac_rows;      // rows per Jacobian tile / partial slot: GROUP_ROWS (128), or REFINE_ROWS (64) in the exact mode
    int jgroups_max;   // jac_rows-row groups per object in the Jacobian kernel grid / partial slots per object
    int fgroups_max;   // 128-row groups per object in the forward (ray-sample) kernel grid
    int n_phases;      // front-to-back ray phases per iteration (1: every in-ball sample in one pass)
    int phase;         // phase the sample / forward launches of this view work on
    // inputs
    const int* pts_off;   // [B+1]
    const int* ray_off;   // [B+1]
    const int* n_fg;      // [B]
    const float* pts;     // [total_pts][3]  camera frame
    const float* rays;    // [total_rays][3]
    const float* depth;   // [total_rays]
    // state
    ObjState* st;         // [B]
    float* code;          // [B][MAX_L]
    float* c0;            // [B][512] latent bias of layer 0
    float* c4;            // [B][512] latent bias of layer 4
    // scratch
    float4* frows;        // [total_rays*D]  compacted in-ball samples per object: x,y,z,dest; after the forward
                          //                 phases: staging of the selected render rows, ray r at [r*D, r*D+cnt)
    float* sdf_samp;      // [total_rays*D]  SDF per ray sample (SDF_OUTSIDE outside the ball)
    float4* rrows;        // [total_rays*D]  per object: samples the FP16 forward pass left within refine_eps of a decision
                          //                 threshold (| |s| - th | < eps): x,y,z,dest; re-evaluated by k_tc_jacx<2>
    float refine_eps;     // 0: no second evaluation (FP32 mode, or switched off)
    int jac_wsplit;       // 1: the tcgen05 Jacobian kernel multiplies by hi(W) + lo(W) (two MMA passes per K chunk)
    int* ray_cnt;         // [total_rays]
    int* ray_roff;        // [total_rays]    exclusive row offset of the ray inside its 128-ray chunk
    int* chunk_cnt;       // [B][ray_chunks] rows selected by each chunk of rays, then (scanned) the chunk's first row
    int ray_chunks;       // 128-ray chunks per object in the k_gn_ray_count grid (from max_rays)
    float* ray_res;       // [total_rays]    depth residual of the ray (shared by all its selected rows)
    int* ray_hit;         // [total_rays]    first sample with sdf <= -th in the previous iteration (-1: none yet, D: no hit)
    float4* jrows;        // [total_pts + total_rays*D]  per object: M surface rows then K render rows: x,y,z,coef
    float* jres;          // same indexing: residual of render rows
    uint8_t* active;      // [total_pts]     pose-only inlier flags
    float* sres;          // [total_pts]     last SDF residual per surface point
    float* partials;      // [B][jgroups_max][PART_STRIDE]
    double* hred;         // [B][PART_STRIDE] the partials summed in tile order (k_gn_reduce -> k_gn_solve)
    // tcgen05 path work lists (filled by k_gn_plan_*)
    int* tile_obj;        // [max tiles] object of each 128-row tile
    int* tile_row0;       // [max tiles] first row of the tile inside the object
    int* n_tiles;         // [4] number of forward / Jacobian / refinement / early Jacobian tiles
    int early;            // 1: the Jacobian tiles made of surface rows only have their own list (etile_*, n_tiles[3]),
                          //    planned once, and run on the handle's second stream beside the ray pipeline
    int* etile_obj;       // [B * ceil(max_pts / jac_rows)]
    int* etile_row0;
    int* tile_off;        // [B+1] first tile of each object in the list being planned (scratch of plan_tiles)
    unsigned int* done;   // [2] "last CTA done" tickets of k_gn_sample / k_gn_ray_count (zero between launches)
    unsigned int* done_obj; // [B] per-object tickets of k_gn_ray_count
    float* dtab;          // [B][64] sampled ray depths of the iteration (linspace(dmin, dmax, D), optimizer.py:126)
    int tc;               // 1: tcgen05 mode (extras images and tile lists are maintained)
    uint8_t* tc_xtra;     // [B][2][16 KB] per-object extra blocks of layers 0 and 4 (kernels_tc.cu)
    unsigned long long* stats;  // null or [2]: algorithmic rows evaluated (forward-only, SDF+Jacobian)
};

cudaError_t simt_configure();
void launch_latent_bias(const DecoderDev& dec, const float* codes, int n_codes, int stride, float* c0,
                        float* c4, cudaStream_t st);
void launch_sdf_rows_simt(const DecoderDev& dec, const float* xyz, int64_t n, const float* c0,
                          const float* c4, float* out, cudaStream_t st);
void launch_grid_simt(const DecoderDev& dec, int G, int64_t i0, int64_t i1, int coord_mode, const float* c0,
                      const float* c4, float* out, cudaStream_t st);
void launch_sdf_jac_simt(const DecoderDev& dec, const float* xyz, int64_t n, const float* c0, const float* c4,
                         float* sdf_out, float* jac_out, cudaStream_t st);
void launch_mean(const float* v, int64_t n, float* out, cudaStream_t st);

void launch_gn_init(const GnWork& w, const float* pose_in, const float* code_in, int code_stride,
                    const float* scale_in, cudaStream_t st);
void launch_gn_prep(const GnWork& w, int iter, cudaStream_t st);
void launch_gn_final(const GnWork& w, float* pose_out, float* code_out, int code_stride, const float* scale_in,
                     int* status_out, float* loss_out, cudaStream_t st);
void launch_gn_sample(const GnWork& w, cudaStream_t st);
void launch_gn_fwd_simt(const GnWork& w, cudaStream_t st);
void launch_gn_refine_plan(const GnWork& w, cudaStream_t st);
void launch_gn_rays(const GnWork& w, int max_rays, cudaStream_t st);
void launch_gn_jac_simt(const GnWork& w, cudaStream_t st);
void launch_gn_solve(const GnWork& w, int iter, int last_iter, float* dbg, int dbg_stride, cudaStream_t st);


// marching cubes (mesh.cu)
cudaError_t mc_configure();
size_t mc_scratch_bytes(int G);
int launch_marching_cubes(const float* sdf, int G, float level, float spacing, float origin, int max_verts,
                          int max_tris, float* verts, int* faces, int* counts, void* scratch, cudaStream_t st);

}  // namespace dsp

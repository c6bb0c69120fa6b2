// tcgen05 (FP16 operands, FP32 TMEM accumulators) decoder kernels: interface to api.cu.
#pragma once
#include "kernels.h"

namespace dsp {

// Packs the FP16 weight stage images and the stage programs, uploads them (device pointer in *hblob,
// also recorded in dec) and sets the kernels' shared-memory attributes.
// Returns 0 on success, 1 if the decoder geometry is unsupported by this path, 2 on a CUDA error.
int tc_pack_and_configure(DecoderDev& dec, const float* const* W_host, const float* const* b_host,
                          __half** hblob);
const char* tc_last_error();
size_t tc_xtra_bytes_per_object();

// src_mode 0: xyz array; 1: voxel grid, reference shear; 2: voxel grid, lattice.
// All launchers return the number of kernels launched.
int tc_launch_sdf_rows(const DecoderDev& dec, int sm_count, const float* xyz, int64_t n, int64_t i0, int G,
                       int src_mode, const float* c0, const float* c4, float* out, cudaStream_t st);
int tc_launch_sdf_jac(const DecoderDev& dec, int sm_count, const float* xyz, int64_t n, const float* c0,
                      const float* c4, float* sdf_out, float* jac_out, int wsplit, cudaStream_t st);
int tc_launch_gn_fwd(const GnWork& w, int sm_count, cudaStream_t st);
int tc_launch_gn_refine(const GnWork& w, int sm_count, cudaStream_t st);
// FP32-grade SDF + Jacobian on the tensor cores (hi/lo operand pairs, 64-row tiles): DSPSLAM_MODE_TC_EXACT
int tc_launch_sdf_jac_exact(const DecoderDev& dec, int sm_count, const float* xyz, int64_t n, const float* c0,
                            const float* c4, float* sdf_out, float* jac_out, cudaStream_t st);
int tc_launch_gn_jac_exact(const GnWork& w, int sm_count, cudaStream_t st);
int tc_launch_gn_jac(const GnWork& w, int sm_count, int early, cudaStream_t st);

}  // namespace dsp

// placeholder until the tcgen05 kernels land (next commit): TC mode reports "unavailable"
#include "kernels_tc.h"
namespace dsp {
int tc_pack_and_configure(DecoderDev&, const float* const*, const float* const*, __half**) { return 1; }
const char* tc_last_error() { return "tcgen05 kernels not built"; }
void tc_launch_sdf_rows(const DecoderDev&, int, const float*, int64_t, int64_t, int, int, const float*, const float*, float*, cudaStream_t) {}
void tc_launch_sdf_jac(const DecoderDev&, int, const float*, int64_t, const float*, const float*, float*, float*, cudaStream_t) {}
int tc_launch_gn_fwd(const GnWork&, int, cudaStream_t) { return 0; }
int tc_launch_gn_jac(const GnWork&, int, cudaStream_t) { return 0; }
}

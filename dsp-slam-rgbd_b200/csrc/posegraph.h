// Launchers of posegraph.cu (camera-object pose-graph edges).  Each returns the
// number of kernels it launched.  Poses are g2o Vector7d rows (tx, ty, tz, qx, qy, qz, qw), FP64, on the device.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dsp {

int launch_pg_from_matrices(const float* T, int n, double* out7, cudaStream_t st);
// A/b (per vertex) and Hij (per edge) are optional (nullptr skips the quadratic form)
int launch_pg_linearize(const double* verts, int nv, const int32_t* ei, const int32_t* ej, const double* meas, int ne,
                        double inv_sigma, double huber_delta, double* err, double* Ji, double* Jj, double* chi2,
                        double* rho, double* wgt, double* A, double* b, double* Hij, cudaStream_t st);
int launch_pg_oplus(double* verts, const double* upd, int n, cudaStream_t st);

}  // namespace dsp

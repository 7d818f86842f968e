// kernels.h -- internal launch functions (one per .cu); all enqueue on `stream` and return SHOPT_* codes.
#pragma once
#include <cuda_runtime.h>

namespace shopt {

// lbfgs_fit.cu
int fit_analytic_launch(const double *d_stamps, int r, int c, long long N, int n_a, const double *d_x0, double g_tol,
                        int max_iter, double *d_s, double *d_g1, double *d_g2, int *d_status, int *d_iters,
                        unsigned long long *d_counter, cudaStream_t stream, double *d_f = nullptr,
                        double *d_g = nullptr);   // d_f/d_g non-null: evaluate cost and gradient at x0 only

}  // namespace shopt

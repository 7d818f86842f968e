// kernels.h -- internal launch functions (one per .cu); all enqueue on `stream` and return SHOPT_* codes.
#pragma once
#include <cuda_runtime.h>
#include <vector>

namespace shopt {

// api_field.cu (host): min |A x - b_j|, A n x 10 column-major, B n x nrhs; both overwritten; returns the rank
int field_lstsq(std::vector<double> &A, std::vector<double> &B, long long n, int nrhs, double *coef);
// the same from the normal equations (G 10 x 10, b [nrhs][10]); pivoted Cholesky; returns the rank
int field_solve_normal(const double *G, const double *b, int nrhs, double *coef);

// lbfgs_fit.cu
int fit_analytic_launch(const double *d_stamps, int r, int c, long long N, int n_a, const double *d_x0, double g_tol,
                        int max_iter, double *d_s, double *d_g1, double *d_g2, int *d_status, int *d_iters,
                        unsigned long long *d_counter, cudaStream_t stream, double *d_f = nullptr,
                        double *d_g = nullptr);   // d_f/d_g non-null: evaluate cost and gradient at x0 only

// polyfit.cu  (rc = r*c; idx may be null = stars 0..n-1; scratch: polyfit_scratch_bytes(rc, n, degree))
size_t polyfit_scratch_bytes(long long rc, long long n, int degree);
int polyfit_normal_launch(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                          long long n, int degree, double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st);
int polyfit_solve_launch(const double *acc, long long rc, int degree, double *poly, int *info, cudaStream_t st);
int polyfit_mse_launch(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                       long long n, int degree, const double *poly, double *mse, double *scratch, size_t scratch_bytes,
                       cudaStream_t st);
// per-pixel median over the listed stars: mode 1 -> med[rc] (median_constraint); mode 2 -> re-accumulate V'P in `acc`
// for the pixels whose sums are NaN, with NaN values replaced by the median (interpolate.jl:126-130)
int polyfit_median_launch(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                          long long n, int degree, int mode, double *acc, double *med, cudaStream_t st);
int polyfit_set_coef0_launch(const double *med, long long rc, double *poly, cudaStream_t st);
int psf_at_launch(const double *poly, long long rc, int degree, const double *u, const double *v, long long M,
                  double *out, cudaStream_t st);
int small_median_launch(const double *pgf, long long rc, const long long *idx, int n, int degree, double *poly,
                        cudaStream_t st);

// pixelgrid.cu  (scratch: pca_scratch_bytes(r, c, ncomp), 0 when the stamp fits shared memory)
size_t pca_scratch_bytes(int r, int c, int ncomp);
int pca_launch(const double *d_stamps, int r, int c, long long N, int ncomp, double pratio, int lanczos_a, int unity_sum,
               double *d_pgf, int *d_k, int *d_status, unsigned long long *d_counter, double *scratch,
               size_t scratch_bytes, cudaStream_t st);
// element-wise pixel-grid fits: smooth mode with unity_sum (nanToZero ./ sum) and chisq mode (closed-form minimiser)
int elem_pgf_launch(const double *d_in, const double *d_w, const double *d_x0, double fill, long long rc, long long N,
                    int chisq, int unity, double *d_out, cudaStream_t st);
int smooth_launch(const double *d_in, int r, int c, long long N, int lanczos_a, int nan_to_zero_flag, double *d_out,
                  cudaStream_t st);

}  // namespace shopt

// kernels.h -- internal launch functions (one per .cu); all enqueue on `stream` and return SHOPT_* codes.
#pragma once
#include <cuda_runtime.h>
#include <vector>

namespace shopt {

// lbfgs_fit.cu
int fit_analytic_launch(const double *d_stamps, int r, int c, long long N, int n_a, const double *d_x0, double g_tol,
                        int max_iter, double *d_s, double *d_g1, double *d_g2, int *d_status, int *d_iters,
                        unsigned long long *d_counter, cudaStream_t stream, double *d_f = nullptr,
                        double *d_g = nullptr,    // d_f/d_g non-null: evaluate cost and gradient at x0 only
                        int profile = 0);         // SHOPT_PROFILE_*

// polyfit.cu  (rc = r*c; scratch: polyfit_scratch_bytes(rc, n, degree))
// A list of stars: idx may be null = stars 0..n-1.  When n_dev is set the ACTUAL length lives on the device and n is
// only an upper bound used to size grids and scratch (the stage's training list shrinks without the host knowing).
struct StarList {
    const long long *idx = nullptr;
    long long n = 0;
    const long long *n_dev = nullptr;
};
struct PolyFactor;
struct Exchange;
size_t polyfit_scratch_bytes(long long rc, long long n, int degree);
size_t polyfit_mse_part_len(long long rc, long long n, int degree);   // doubles of scratch of polyfit_mse_launch alone
// ctr: device {u0, 1/su, v0, 1/sv} of the centred basis, or null = raw monomials (polyfit_factor.h)
int polyfit_normal_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                          const double *ctr, double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st);
int polyfit_factor_launch(const double *G, int degree, const double *ctr, const long long *nrows_dev, long long nrows,
                          int drop_const, PolyFactor *F, cudaStream_t st);
int polyfit_solve_launch(const double *acc, long long rc, int degree, const PolyFactor *F, const PolyFactor *Fred,
                         const double *med, double *poly, cudaStream_t st);
int polyfit_mse_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                       const double *poly, double *mse, double *scratch, size_t scratch_bytes, cudaStream_t st);
// per-pixel median over the listed stars: mode 1 -> med[rc] (median_constraint); mode 2 -> re-accumulate V'P in `acc`
// for the pixels whose sums are NaN, with NaN values replaced by the median (interpolate.jl:126-130)
int polyfit_median_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                          const double *ctr, int mode, double *acc, double *med, cudaStream_t st);
int polyfit_sub_launch(double *acc, const double *sub, long long len, cudaStream_t st);
size_t polyfit_shard_median_bytes(long long rc, int degree);
int polyfit_median_sharded_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                                  const double *ctr, int mode, double *acc, double *med, void *work, const Exchange &ex,
                                  cudaStream_t st);
int psf_at_launch(const double *poly, long long rc, int degree, const double *u, const double *v, long long M,
                  double *out, cudaStream_t st);
int small_median_launch(const double *rows, long long rc, int n, int degree, double *poly, cudaStream_t st);
int small_gather_launch(const double *pgf, long long rc, const long long *idx, const unsigned *pos, const long long *n_dev,
                        long long NG, unsigned *bits, double *rows, const Exchange &ex, cudaStream_t st);

// api_ingest.cu: nanMask / unity sum / crop / resample of raw vignets (float32 or float64) into float64 stamps
int ingest_launch(const void *d_in, int f32, int r0, int c0, long long N, int unity, int new_dim, int lanczos_a,
                  double *d_out, double *d_tmp, cudaStream_t st);

// select.cu  (device-side control flow of the stage; StageState in stage_state.h)
struct StageState;
int flag_launch(const double *s_model, const int *pca_status, const double *s_learned, const int *learned_status, long long n,
                double lo, double hi, int *flags, int *counters, cudaStream_t st);
int surv_bits_launch(const long long *perm, long long NG, long long off, long long n, const int *flags, unsigned *lbits,
                     unsigned *gbits, int *bad_perm, cudaStream_t st);
int split_launch(const unsigned *gbits, long long NG, double ratio, const int *counters, StageState *state, cudaStream_t st);
int training_list_launch(const unsigned *lbits, const long long *perm, long long NG, long long off, StageState *state,
                         long long *out_idx, unsigned *out_pos, int *flags, int *blockcnt, long long *blockoff,
                         cudaStream_t st);
int select_worst_launch(const double *mse, const unsigned *pos, long long n_bound, const long long *n_ptr, long long m,
                        StageState *state, unsigned *hist, const Exchange &ex, cudaStream_t st);
int clip_threshold_launch(const double *mse, const long long *n_ptr, double nc, StageState *state, double *sums,
                          const Exchange &ex, cudaStream_t st);
int compact_training_launch(const double *mse, const unsigned *pos, long long n_bound, int cur, int sigma, StageState *state,
                            const long long *in_idx, long long *out_idx, unsigned *out_pos, int *flags, int *blockcnt,
                            long long *blockoff, cudaStream_t st, long long *rej_idx = nullptr);
int uv_centre_launch(const double *u, const double *v, long long n, double *sums, double *ctr, const Exchange &ex,
                     cudaStream_t st);

// pixelgrid.cu  (scratch: pca_scratch_bytes(r, c, ncomp), 0 when the stamp fits shared memory)
size_t pca_scratch_bytes(int r, int c, int ncomp, long long N);
// pca_power.cu: stage 1 of the PCA pixel-grid fit (stars whose reconstruction has one component)
struct PcaPowerArgs {
    const double *stamps;
    int r, c;
    long long N;
    double pratio;
    int lanczos_a, unity_sum;
    double *pgf;
    int *k_out, *status;
    unsigned long long *counter;        // work counter (zeroed by the launch)
    long long *defer_list;              // stars handed over to pca_kernel ...
    unsigned long long *defer_count;    // ... and how many (zeroed by the launch)
    double kz[15];                      // 1-D lanczos factor (2a+1 taps)
};
int pca_power_launch(const PcaPowerArgs &a, cudaStream_t st);
int pca_launch(const double *d_stamps, int r, int c, long long N, int ncomp, double pratio, int lanczos_a, int unity_sum,
               double *d_pgf, int *d_k, int *d_status, unsigned long long *d_counter, double *scratch,
               size_t scratch_bytes, cudaStream_t st);
// element-wise pixel-grid fits: smooth mode with unity_sum (nanToZero ./ sum) and chisq mode (closed-form minimiser)
int elem_pgf_launch(const double *d_in, const double *d_w, const double *d_x0, double fill, long long rc, long long N,
                    int chisq, int unity, double *d_out, cudaStream_t st);
int smooth_launch(const double *d_in, int r, int c, long long N, int lanczos_a, int nan_to_zero_flag, double *d_out,
                  cudaStream_t st);

}  // namespace shopt

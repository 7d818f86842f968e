/*
 * shopt_b200.h -- C ABI of libshopt_b200.so: the per-star PSF fitting stage of ShOpt.jl on B200.
 *
 * The reference (EdwardBerman/shopt) has no FFI/plugin interface: the stage is a set of script-level
 * loops in shopt/shopt.jl.  Each entry point below replaces one of those loops with ONE batched call
 * and is what a thin Julia `ccall` shim binds (INTEGRATION.md shows the shim).  Paths cited are
 * relative to the reference checkout.
 *
 * Conventions
 *   - FP64 throughout.  Julia column-major: a batch of N stamps of r x c pixels is the Julia array
 *     Array{Float64,3}(r, c, N); pixel (u,v) (1-based) of star i (0-based) is stamps[i*r*c + (u-1) + (v-1)*r].
 *   - All indices crossing this interface are 0-based (the Julia shim adds/subtracts 1).
 *   - Host entry points (no suffix) take HOST pointers, copy H2D / D2H themselves and block until done.
 *     `_dev` entry points take DEVICE pointers on the current device, enqueue on `stream`
 *     (a cudaStream_t passed as void*) and do not synchronise; they are what the multi-GPU host code
 *     and the benchmark's device-resident leg call.
 *   - Every function returns SHOPT_OK (0) or a negative SHOPT_E_* code; nothing throws or aborts.
 *     Per-star outcomes are reported in `status` arrays (the reference's per-star try/catch).
 *   - There is no CPU fallback: without a CUDA device every compute call returns SHOPT_E_NODEVICE.
 */
#ifndef SHOPT_B200_H
#define SHOPT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SHOPT_OK              0
#define SHOPT_E_NODEVICE     -1   /* no CUDA device / driver                                   */
#define SHOPT_E_BADARG       -2   /* null pointer, non-positive size, n_a > min(r,c), degree>8  */
#define SHOPT_E_CUDA         -3   /* a CUDA runtime call failed; see shopt_last_error()         */
#define SHOPT_E_NOMEM        -4   /* device allocation failed                                   */
#define SHOPT_E_UNSUPPORTED  -5   /* size outside what the kernels are built for                */

/* per-star status of one analytic fit (same meaning as the oracle's ORC_FIT_*) */
#define SHOPT_FIT_CONVERGED   0   /* Optim assess_convergence: x-, f- or g-convergence          */
#define SHOPT_FIT_MAXITER     1   /* iteration cap reached; minimiser = last iterate            */
#define SHOPT_FIT_LS_FAILED   2   /* LineSearchException: step of the exception taken, stop     */
#define SHOPT_FIT_NONFINITE   3   /* non-finite value/gradient: stop, keep the current x        */
#define SHOPT_FIT_THREW      -1   /* an assert inside the line search fired -> shopt.jl:155-162
                                     catch: s = g1 = g2 = 0                                      */

/* per-star status of one pixel-grid (PCA) fit */
#define SHOPT_PCA_OK          0
#define SHOPT_PCA_FAILED      1   /* non-finite input -> shopt.jl:305-309: zeros(r,c), failedStars */
#define SHOPT_PCA_NOCONV      2   /* subspace iteration hit its cap; result is the last iterate */

/* ---- library ---------------------------------------------------------------------------- */
int         shopt_b200_version(void);          /* 10000*major + 100*minor + patch               */
int         shopt_device_count(void);          /* 0 when no device/driver                       */
int         shopt_set_device(int dev);
const char *shopt_last_error(void);            /* text of the last CUDA error on this thread    */
/* wall-clock-free device timing of the last host call, milliseconds, per stage:
 * [0]=H2D [1]=kernels [2]=D2H (CUDA events on the library's stream) */
int         shopt_last_timing(double ms[3]);
/* The host entry points keep their large device blocks (stamp cube, pixel-grid cube, scratch) between calls on the
 * calling thread's context; this frees them. */
int         shopt_release_buffers(void);
/* kernels this process has launched so far (every launch function of the library counts what it enqueues) */
int64_t     shopt_launch_count(void);
/* measured FP64 FMA throughput of the current device in TFLOP/s (a 10 ms micro-benchmark; the compute roofline of the
 * FP64-bound kernels, reported by bench.py next to the HBM roofline) */
int         shopt_bench_dfma(double *tflops_out);

/* ---- communicator of a sharded run (one process per GPU) --------------------------------------------
 * The reference has no distributed mode; stars shard naturally and the stage's only exchanges are sum-all-reduces of
 * small device buffers.  The library owns its NCCL communicator (bound at run time with dlopen, so single-GPU users
 * need no NCCL): rank 0 calls shopt_comm_unique_id and hands the 128 bytes to the other ranks by any means (a file,
 * MPI, torch.distributed, Julia's Distributed); every rank then calls shopt_comm_init on its device.
 * shopt_comm_allreduce is a shopt_allreduce_fn whose `user` is the shopt_comm*: pass the pair to shopt_stage_run_dev /
 * shopt_stage_run_comm and every exchange is enqueued on the stage's stream from C. */
#define SHOPT_UNIQUE_ID_BYTES 128
typedef struct shopt_comm shopt_comm;
int shopt_comm_unique_id(void *id_out /* SHOPT_UNIQUE_ID_BYTES */);
int shopt_comm_init(const void *id_in, int rank, int world, shopt_comm **out);
int shopt_comm_destroy(shopt_comm *c);
int shopt_comm_rank(const shopt_comm *c);
int shopt_comm_world(const shopt_comm *c);
int shopt_comm_allreduce(void *d_buf, int64_t count, int dtype, void *stream, void *user);

/* ---- (0) ingest: the per-star work of cataloging() that precedes the stage --------------------------
 * Replaces, in dataPreprocessing.jl:342-397 (FITS decoding stays with astropy / CFITSIO):
 *   nanMask (masks.jl:5-17: x < -1000 -> NaN);  unity_sum: ./ sum(nanToZero(nanMask(.))) (:344-346);
 *   new_dim < r0: get_middle_nxn crop (:207-216, :384);  new_dim > r0: bilinear oversample_image (:141-162, :391);
 *   a resized stamp is ALWAYS renormalised, nm(.) / sum(nz(nm(.))), and smoothed, smooth(., lanczos_a) (quirk B.9).
 * vignets (r0, c0, N): the raw VIGNET cube, float32 when dtype_f32 != 0 (what SExtractor writes) else float64;
 * stamps_out (new_dim, new_dim, N) float64 (new_dim == r0: (r0, c0, N), no resize).  Resizing needs r0 == c0. */
int shopt_ingest_batch(const void *vignets, int dtype_f32, int r0, int c0, int64_t N, int unity_sum, int new_dim,
                       int lanczos_a, double *stamps_out);
int shopt_ingest_batch_dev(const void *d_vignets, int dtype_f32, int r0, int c0, int64_t N, int unity_sum, int new_dim,
                           int lanczos_a, double *d_stamps_out, void *stream);
/* outliers_filter (outliers.jl:16-34) on the Float32 SNR_WIN column: q1 = quantile(snr, k) (SnRPercentile), returns
 * the count of stars with snr < q1 and their 0-based indices (ascending); SHOPT_E_BADARG when snr holds a NaN
 * (Statistics.quantile raises).  Host function.  q1_out may be NULL. */
int64_t shopt_snr_cut(const float *snr, int64_t N, double k, int64_t *idx_out, double *q1_out);

/* ---- (1) analytic profile fit ------------------------------------------------------------
 * Replaces the loops  shopt/shopt.jl:130-164  (model stars, cost/g!)  and  :357-397  (learned
 * stars, costD/gD!):  Optim.optimize(cost, g!, x0, LBFGS(), Options(g_tol))  per star, with
 * cost = shopt/analyticLBFGS.jl:8-46 on the get_middle_nxn(n_a) crop (dataPreprocessing.jl:207-216),
 * fGaussian = shopt/radialProfiles.jl:5-9, then s = x1^2, (g1,g2) from (e1,e2) (shopt.jl:145-153).
 *   stamps  (r,c,N)      x0 (3,N): the reference draws rand(3) per star -- an explicit input here
 *   g_tol   minAnalyticGradientModel / ...Learned (shopt.yml AnalyticFitParams)
 *   max_iter  Optim default 1000
 *   s,g1,g2 (N) out;  status,iters (N) out, may be NULL                                          */
int shopt_fit_analytic_batch(const double *stamps, int r, int c, int64_t N, int n_a,
                             const double *x0, double g_tol, int max_iter,
                             double *s, double *g1, double *g2, int32_t *status, int32_t *iters);
int shopt_fit_analytic_batch_dev(const double *d_stamps, int r, int c, int64_t N, int n_a,
                                 const double *d_x0, double g_tol, int max_iter,
                                 double *d_s, double *d_g1, double *d_g2,
                                 int32_t *d_status, int32_t *d_iters, void *stream);

/* The same fit with the radial profile as an argument.  SHOPT_PROFILE_GAUSSIAN is fGaussian (radialProfiles.jl:5-9),
 * the only profile the reference's cost ever calls; SHOPT_PROFILE_KOLMOGOROV is fkolmogorov (radialProfiles.jl:11-24,
 * defined but unreachable in the reference) in the same cost: sum-normalised over all crop pixels, residuals over the
 * non-NaN ones.  Fkolmogorov(k, 1) as written is exp(-C k), C = (24 gamma(6.5) 0.2)^(5/6) / (2 pi)^(5/3) = 19.3531...,
 * so the quadgk integral of :21-23 has the closed form (1 / 2 pi) C / (C^2 + squareRadius)^(3/2), which is what the
 * kernel evaluates (no quadrature).  n_a <= 160 for either profile (SHOPT_E_UNSUPPORTED beyond: the crop of one star,
 * and of two when they fit, is staged in one SM's shared memory). */
#define SHOPT_PROFILE_GAUSSIAN   0
#define SHOPT_PROFILE_KOLMOGOROV 1
int shopt_fit_profile_batch(int profile, const double *stamps, int r, int c, int64_t N, int n_a,
                            const double *x0, double g_tol, int max_iter,
                            double *s, double *g1, double *g2, int32_t *status, int32_t *iters);
int shopt_fit_profile_batch_dev(int profile, const double *d_stamps, int r, int c, int64_t N, int n_a,
                                const double *d_x0, double g_tol, int max_iter,
                                double *d_s, double *d_g1, double *d_g2,
                                int32_t *d_status, int32_t *d_iters, void *stream);
int shopt_profile_cost_grad_batch(int profile, const double *stamps, int r, int c, int64_t N, int n_a,
                                  const double *x, double *f, double *g);
int shopt_profile_cost_grad_batch_dev(int profile, const double *d_stamps, int r, int c, int64_t N, int n_a,
                                      const double *d_x, double *d_f, double *d_g, void *stream);

/* cost(params) / g!(storage, params) of analyticLBFGS.jl:8-46, 89-94 (costD / gD! :49-87, 96-101 on
 * pixel-grid fits) for one parameter vector [sigma, e1, e2] per star:  f (N), g (3,N). */
int shopt_analytic_cost_grad_batch(const double *stamps, int r, int c, int64_t N, int n_a,
                                   const double *x, double *f, double *g);
int shopt_analytic_cost_grad_batch_dev(const double *d_stamps, int r, int c, int64_t N, int n_a,
                                       const double *d_x, double *d_f, double *d_g, void *stream);

/* s-blacklist (shopt.jl:166-171) and learned-fit failed list (shopt.jl:384-386):
 * 0-based indices with s < s_lo or s > s_hi, ascending; returns the count (host arrays). */
int64_t shopt_s_out_of_bounds(const double *s, int64_t N, double s_lo, double s_hi, int64_t *idx);

/* ---- (2) pixel-grid fit, mode == "PCA" ------------------------------------------------------
 * Replaces the loop shopt/shopt.jl:281-314: nanToZero (masks.jl:38-50, NaN -> 0.001), pca_image
 * (pca.jl:1-18: MultivariateStats fit(PCA; maxoutdim=ncomp, pratio) / transform / reconstruct with
 * rows = variables, columns = observations), then smooth(., lanczos_a)
 * (dataPreprocessing.jl:94-139) when unity_sum == 0, or division by the stamp sum WITHOUT smoothing
 * when unity_sum != 0 (the reference's 1-argument smooth() call at shopt.jl:292 always raises, so
 * its catch branch :295 runs).  lanczos_a <= 0 skips the smoothing.
 *   pgf (r,c,N) out;  k_out (N): number of components kept;  status (N): SHOPT_PCA_*             */
int shopt_pca_batch(const double *stamps, int r, int c, int64_t N, int ncomp, double pratio,
                    int lanczos_a, int unity_sum, double *pgf, int32_t *k_out, int32_t *status);
int shopt_pca_batch_dev(const double *d_stamps, int r, int c, int64_t N, int ncomp, double pratio,
                        int lanczos_a, int unity_sum, double *d_pgf, int32_t *d_k_out,
                        int32_t *d_status, void *stream);

/* smooth(data, a) alone (dataPreprocessing.jl:112-139; mode == "smooth", shopt.jl:316-348, applies
 * it to nanToZero(star) -- pass nan_to_zero = 1 for that). */
int shopt_smooth_batch(const double *stamps, int r, int c, int64_t N, int lanczos_a, int nan_to_zero,
                       double *out);
int shopt_smooth_batch_dev(const double *d_stamps, int r, int c, int64_t N, int lanczos_a,
                           int nan_to_zero, double *d_out, void *stream);

/* The whole loop body of mode == "smooth" (shopt.jl:316-348) per star: data = nanToZero(star); unity_sum == 0 ->
 * smooth(data, lanczos_a); unity_sum != 0 -> data ./ sum(data) (the reference's smooth(data) call with one argument
 * raises, so its catch branch :330 runs and nothing is smoothed). */
int shopt_smooth_mode_batch(const double *stamps, int r, int c, int64_t N, int lanczos_a, int unity_sum, double *pgf);
int shopt_smooth_mode_batch_dev(const double *d_stamps, int r, int c, int64_t N, int lanczos_a, int unity_sum,
                                double *d_pgf, void *stream);

/* mode == "chisq" (shopt.jl:194-224 with chisq_cost / chisq_g!, chisq.jl:1-11): the reference minimises
 *   sum_px isfinite(t) ? t : 0,  t = (p - star)^2 / weight     over the r*c pixel values p by LBFGS from rand(r*c).
 * The objective is separable and quadratic per pixel, so the minimiser is written down directly: p = star wherever the
 * term is a finite convex quadratic (star finite, weight finite and > 0); every other pixel has zero gradient and
 * keeps its START value -- x0 (r,c,N), the reference's rand(r*c), an explicit input; x0 == NULL -> `fill`.
 * (LBFGS stops at |2 (p - star) / weight| <= g_tol, i.e. within g_tol * weight / 2 of this answer.)
 * unity_sum != 0 divides each fit by its sum (shopt.jl:216-218).  weights (r,c,N) = errVignets. */
int shopt_chisq_batch(const double *stamps, const double *weights, const double *x0, double fill, int r, int c,
                      int64_t N, int unity_sum, double *pgf);
int shopt_chisq_batch_dev(const double *d_stamps, const double *d_weights, const double *d_x0, double fill, int r, int c,
                          int64_t N, int unity_sum, double *d_pgf, void *stream);

/* ---- (3) per-pixel polynomial interpolation with outlier-star rejection ----------------------
 * Replaces shopt/shopt.jl:492-577: `iterations` rounds of
 *   polynomial_optimizer(degree, u, v, z) per pixel (interpolate.jl:105-147; term order p outer,
 *   q inner, A[i,k] = u_i^p v_i^q, least squares),
 *   per-star MSE of the reconstruction (shopt.jl:496-511, interpolate.jl:89-102),
 *   worst_10_percent (shopt.jl:513-520; sigma_clip != 0 -> the mean+3sigma variant :522-532),
 *   rejected stars dropped except after the last round (shopt.jl:557-575).
 * The least squares is solved through the normal equations V'V c = V'z (one K x K Cholesky shared
 * by all pixels); median_constraint (interpolate.jl:118-124,139-141) and fewer than 5 training
 * stars (:142-145) follow the reference's overrides.
 *   pgf (r,c,N); u,v (N); train_idx (n_train) 0-based, the randperm split of shopt.jl:472-490 is an
 *   INPUT;  poly_out (r,c,K) K=(degree+1)(degree+2)/2;  mse_out (>= n_train): last round's training
 *   errors in that round's order;  kept_idx/n_kept: training set after the last applied rejection;
 *   n_per_round (iterations): training-set size of each round.  Outputs other than poly_out may be NULL. */
int shopt_polyfit(const double *pgf, int r, int c, int64_t N, const double *u, const double *v,
                  const int64_t *train_idx, int64_t n_train, int degree, int iterations,
                  int median_constraint, int sigma_clip, double *poly_out, double *mse_out,
                  int64_t *kept_idx, int64_t *n_kept, int64_t *n_per_round);

/* Building blocks of one round on device buffers (what shopt_polyfit and the stage call).
 *  normal_dev : acc[0..K*K) = V'V, acc[K*K .. K*K + K*r*c) = V'P over the listed stars of this shard (layout
 *               V'P[k*rc + px]); the buffer every rank all-reduces.  V is built from CENTRED, SCALED coordinates
 *               ((u - u0)/su, (v - v0)/sv) when d_centre = {u0, 1/su, v0, 1/sv} (device) is given, raw when NULL:
 *               the reference passes raw sky coordinates (RA ~ 150, Dec ~ 2), for which V'V in raw monomials is useless.
 *  solve_dev  : A \ z of interpolate.jl:138 for every pixel from those normal equations: pivoted Cholesky in the
 *               centred basis, then the K x K change of basis is factored by column-pivoted QR with LAPACK's rank rule
 *               (rcond = min(n,K) eps) and minimum-norm completion, so poly (r,c,K) holds RAW-monomial coefficients with
 *               the reference's meaning whether the design is well conditioned, ill conditioned or rank deficient.
 *               n_rows = rows of the design matrix (training stars).  d_info (2 ints, device): numerical rank, SHOPT_POLY_* bits.
 *  mse_dev    : mse[t] = mean_px (sum_k poly[px,k] u_t^p v_t^q - pgf[idx_t, px])^2 for the listed stars (raw u, v). */
#define SHOPT_POLY_RANK_DEFICIENT  1  /* raw design numerically rank deficient: minimum-norm answer, like A \ z     */
#define SHOPT_POLY_NONFINITE       2  /* NaN / Inf in the normal equations: coefficients are NaN                    */
#define SHOPT_POLY_ILL_CONDITIONED 4  /* centred design nearly dependent: the normal equations lose digits          */
int64_t shopt_polyfit_acc_len(int r, int c, int degree);                   /* K*K + K*r*c doubles */
int shopt_polyfit_centre_dev(const double *d_u, const double *d_v, int64_t n, double *d_centre /* 4 */, void *stream);
int shopt_polyfit_normal_dev(const double *d_pgf, int r, int c, const double *d_u, const double *d_v,
                             const int64_t *d_idx, int64_t n, int degree, const double *d_centre, double *d_acc,
                             void *stream);
int shopt_polyfit_solve_dev(const double *d_acc, int r, int c, int degree, const double *d_centre, int64_t n_rows,
                            double *d_poly, int32_t *d_info, void *stream);
/* The K x K part of that solve on the host (no device needed): G = V'V (K x K, row-major) in the basis `centre`
 * describes (NULL = raw), b (K, nrhs) right-hand sides V'z stored b[k * nrhs + j] -> coef[k * nrhs + j] raw-monomial
 * coefficients.  Returns the numerical rank (K = full) or a negative SHOPT_E_*; info_out receives SHOPT_POLY_* bits. */
int shopt_polyfit_factor_host(const double *G, int degree, const double *centre, int64_t n_rows, const double *b,
                              int nrhs, double *coef, int32_t *info_out);
int shopt_polyfit_mse_dev(const double *d_pgf, int r, int c, const double *d_u, const double *d_v,
                          const int64_t *d_idx, int64_t n, int degree, const double *d_poly,
                          double *d_mse, void *stream);
/* shopt_polyfit_mse_dev for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels and bit-identical results, but the per-tile partial sums live in a static
 * __device__ array of SHOPT_NOALLOC_MSE_SCRATCH_BYTES (held by every GPU that loads the library until exit) and the call
 * performs no cudaMalloc / cudaFree and creates no library context.  A list of more than
 * shopt_polyfit_mse_noalloc_max_stars(r, c, degree) stars is rejected with SHOPT_E_UNSUPPORTED and a text in
 * shopt_last_error(); 7 710 stars of 131 x 131 at degree <= 3.  All calls share the array: calls that overlap on the GPU
 * (two streams, concurrent graph replays) race and must be serialised by the caller. */
#define SHOPT_NOALLOC_MSE_SCRATCH_BYTES (8 << 20)
int64_t shopt_polyfit_mse_noalloc_max_stars(int r, int c, int degree);
int shopt_polyfit_mse_noalloc_dev(const double *d_pgf, int r, int c, const double *d_u, const double *d_v,
                                  const int64_t *d_idx, int64_t n, int degree, const double *d_poly,
                                  double *d_mse, void *stream);
/* worst_10_percent on host arrays: returns count, idx_out = positions in err (stable descending). */
int64_t shopt_worst_10_percent(const double *err, int64_t n, int sigma_clip, int64_t *idx_out);

/* ---- consumer: reader.jl:40-68  p(u,v,polMatrix,degree) batched over M positions -------------
 * stars_out (r,c,M): polynomial evaluated at (u_m,v_m), each stamp divided by its sum. */
int shopt_psf_at_batch(const double *poly, int r, int c, int degree, const double *u, const double *v,
                       int64_t M, double *stars_out);
int shopt_psf_at_batch_dev(const double *d_poly, int r, int c, int degree, const double *d_u,
                           const double *d_v, int64_t M, double *d_stars_out, void *stream);

/* ---- s(u,v), g1(u,v), g2(u,v) across the field of view --------------------------------------------
 * Replaces interpolate.jl:7-87 and the three optimize(interpCost*, polyG_*!, rand(10), LBFGS()) calls of
 * shopt.jl:422-466.  The cost  sum_i (I_p(u_i, v_i) - t_i)^2  with
 *   I_p(u,v) = p1 u^3 + p2 v^3 + p3 u^2 v + p4 v^2 u + p5 u^2 + p6 v^2 + p7 u v + p8 u + p9 v + p10
 * is a convex quadratic, so LBFGS's limit from any start is the least-squares solution; it is computed directly
 * (Householder QR, host code, one factorisation for all `nvalues` right-hand sides).
 *   values[j] (n): e.g. {s_data, g1_data, g2_data};  coef (10, nvalues): coef[10 j + k], the reference's order.
 * Returns the numerical rank of the design matrix (10 = full; dependent coefficients are set to 0) or SHOPT_E_*. */
int shopt_field_interp(const double *u, const double *v, const double *const *values, int nvalues, int64_t n,
                       double *coef);
/* s(u,v) etc. of shopt.jl:437 at n positions (host). */
int shopt_field_eval(const double *coef, const double *u, const double *v, int64_t n, double *out);

/* ---- summary.shopt: the stage's on-disk product ---------------------------------------------------
 * Writers for the truncated summary file (truncate_summary_file: true, the shipped default; dataOutprocessing.jl:52-84)
 * and the full one (:86-168), and reader for read_shopt of reader.jl:13-38, as a self-contained FITS implementation
 * (astropy is what the reference calls through PyCall):
 *   HDU 0 image POLYNOMIAL_MATRIX  float64, numpy shape (r, c, K);  HDU 1 BINTABLE SUMMARY_STATISTICS, column
 *   POLYNOMIAL_DEGREE;  HDU 2 BINTABLE ANALYTIC_PROFILE, columns s_MATRIX, g1_MATRIX, g2_MATRIX (10 rows each, the
 *   cubic field coefficients in the reference's order).
 * poly is (r, c, K) column-major as everywhere in this interface.  The reader finds the columns by name in any HDU,
 * so the reference's full layout (ANALYTIC_PROFILE = HDU 6) reads too; poly may be NULL to query r, c, degree. */
int shopt_summary_write(const char *path, const double *poly, int r, int c, int degree, const double *sC,
                        const double *g1C, const double *g2C);
/* The full file (truncate_summary_file: false, dataOutprocessing.jl:86-168): seven HDUs -- POLYNOMIAL_MATRIX,
 * SUMMARY_STATISTICS (per-star columns U_COORDINATES ... g2_DATA, the training / validation coordinates and
 * POLYNOMIAL_DEGREE, zero-padded to the longest column as astropy does), VIGNETS (numpy (r, c, N)), PIXEL_GRID_FITS and
 * VALIDATION_STARS (numpy (N, r, c): the reference hands astropy the Vector{Matrix}, not the stacked cube), FLAGS,
 * ANALYTIC_PROFILE.  Cubes are (r, c, n) column-major as everywhere in this interface; flag arrays hold the reference's
 * 1-based star numbers as doubles (it stores them with format 'D').  Arrays of length 0 may be NULL. */
typedef struct shopt_summary_full {
    const double *poly; int r, c, degree;                 /* (r, c, K) PolynomialMatrix                               */
    const double *sC, *g1C, *g2C;                         /* 10 each: the cubic field fits                            */
    int64_t n_stars;                                      /* stars left after the splices                             */
    const double *u, *v, *s_model, *g1_model, *g2_model, *s_data, *g1_data, *g2_data;   /* n_stars each             */
    const double *stars, *pgf;                            /* (r, c, n_stars): starCatalog, pixelGridFits              */
    int64_t n_train; const double *train_u, *train_v;     /* training set after the last applied rejection            */
    int64_t n_valid; const double *valid_u, *valid_v, *valid_stars;   /* validation split; valid_stars (r, c, n_valid) */
    int64_t n_flag_snr, n_flag_blacklist, n_flag_failed;
    const double *flag_snr, *flag_blacklist, *flag_failed;            /* outlier_indices, s_blacklist, failedStars     */
} shopt_summary_full;
int shopt_summary_write_full(const char *path, const shopt_summary_full *f);
int shopt_summary_read(const char *path, double *poly, int64_t poly_capacity, int *r, int *c, int *degree, double *sC,
                       double *g1C, double *g2C);

/* ---- the whole stage: what shopt.jl's loops :130-184, :281-314, :357-417, :472-577 become ---------
 * model fit -> s-blacklist -> PCA pixel-grid fit -> learned fit -> failed list -> training split -> polyfit rounds.
 * Parameters mirror shopt.yml (shopt_params).  Stars keep their ORIGINAL indices: instead of splicing arrays like the
 * reference (shopt.jl:177-184, 405-417) the stage reports per-star flags, and blacklisted / failed stars are excluded
 * from the training set exactly as the splices would. */
typedef struct shopt_params {
    int    n_a;                 /* AnalyticFitParams.analyticFitStampSize                        */
    double g_tol_model;         /* AnalyticFitParams.minGradientAnalyticModel                    */
    double g_tol_learned;       /* AnalyticFitParams.minGradientAnalyticLearned                  */
    int    max_iter;            /* Optim default 1000                                            */
    double s_lo, s_hi;          /* dataProcessing.sLowerBound / sUpperBound                      */
    int    pca_terms;           /* NNparams.PCAterms (mode "PCA")                                */
    double pratio;              /* MultivariateStats default 0.99                                */
    int    lanczos;             /* lanczos                                                       */
    int    unity_sum;           /* sum_pixel_grid_and_inputs_to_unity                            */
    int    degree;              /* polynomialDegree                                              */
    int    iterations;          /* iterations (polyfit rounds)                                   */
    double training_ratio;      /* training_ratio                                                */
    int    median_constraint;   /* median_constraint                                             */
    int    sigma_clip;          /* outlier_sigma_clipping                                        */
    int    mode;                /* mode: SHOPT_MODE_PCA / _SMOOTH / _CHISQ (pixel-grid fit)      */
    double chisq_fill;          /* mode chisq without a start cube: value of the pixels that have no finite term */
    int    profile;             /* SHOPT_PROFILE_* of the model and learned fits (the reference: Gaussian)       */
} shopt_params;
#define SHOPT_MODE_PCA    0     /* shopt.jl:281-314 */
#define SHOPT_MODE_SMOOTH 1     /* shopt.jl:316-348 (the shipped shopt.yml default) */
#define SHOPT_MODE_CHISQ  2     /* shopt.jl:194-224 */
void shopt_default_params(shopt_params *p);     /* README.md:300-330 documented defaults          */

#define SHOPT_FLAG_BLACKLISTED 1   /* model s out of [s_lo, s_hi]            (shopt.jl:166-171)   */
#define SHOPT_FLAG_PCA_FAILED  2   /* pixel-grid fit failed -> zeros         (shopt.jl:305-309)   */
#define SHOPT_FLAG_LEARNED_OOB 4   /* learned s out of bounds                (shopt.jl:384-386); a learned fit that THREW
                                      has s = 0 from the catch (:387-393) but is NOT pushed to failedStars there       */
#define SHOPT_FLAG_TRAINING    8   /* star was in the training split                              */
#define SHOPT_FLAG_REJECTED   16   /* dropped by worst_10_percent in some round (shopt.jl:557-575) */

/* Sum-all-reduce over the ranks of a sharded run, supplied by the host program (NCCL / torch.distributed / MPI):
 * in-place on a DEVICE buffer of `count` elements, dtype 0 = float64, 1 = int32, enqueued on (or ordered after)
 * `stream`; returns 0 on success.  The only exchange steps of the stage are these all-reduces:
 * the [V'V | V'P] normal-equation partial sums, and two zero-padded gathers (survivor flags, per-star errors). */
typedef int (*shopt_allreduce_fn)(void *d_buf, int64_t count, int dtype, void *stream, void *user);

typedef struct shopt_stage_io {
    /* inputs: this rank's shard (device pointers in shopt_stage_run_dev, host pointers in shopt_stage_run) */
    const double *stamps;        /* (r, c, n_local)                                               */
    int r, c;
    int64_t n_local;
    const double *u, *v;         /* (n_local)                                                     */
    const double *x0_model;      /* (3, n_local) start points of the model fit   (shopt.jl:134)   */
    const double *x0_learned;    /* (3, n_local) start points of the learned fit (shopt.jl:362)   */
    int64_t n_global, offset;    /* size of the whole catalogue and this shard's first star       */
    const int64_t *perm;         /* HOST, (n_global): the randperm of shopt.jl:474 over ORIGINAL indices; the training
                                    set is the first round(training_ratio * n_survivors) survivors in this order   */
    /* outputs, shard-sized */
    double *model;               /* 3 x n_local: s[0..n), g1[0..n), g2[0..n) of the model fit      */
    double *learned;             /* 3 x n_local of the learned fit, same layout                   */
    double *pgf;                 /* (r, c, n_local) pixel-grid fits; may be NULL in shopt_stage_run */
    int32_t *flags;              /* (n_local) SHOPT_FLAG_*                                        */
    int32_t *pca_k;              /* (n_local) components kept; may be NULL                        */
    /* outputs, replicated on every rank */
    double *poly;                /* (r, c, K) PolynomialMatrix                                    */
    int64_t *n_per_round;        /* HOST (iterations): training-set size of each round; may be NULL */
    int64_t counts[4];           /* HOST: blacklisted, failed, training stars, kept after the last applied rejection */
    /* inputs of mode chisq only (same residence as stamps); may be NULL otherwise */
    const double *weights;       /* (r, c, n_local) errVignets                                    */
    const double *x0_pixels;     /* (r, c, n_local) start values (the reference's rand(r*c)); NULL -> chisq_fill */
    /* output, replicated: cubic s(u,v), g1(u,v), g2(u,v) over the surviving stars' learned fits (shopt.jl:422-466),
     * field[10 j + k], j = s, g1, g2, the reference's coefficient order (see shopt_field_interp); HOST */
    double field[30];
    int32_t field_rank;          /* numerical rank of the cubic design matrix (10 = full)          */
    /* ---- added in version 2 ---- */
    int32_t polyfit_rank;        /* out: numerical rank of the last round's design matrix (K = full; 0 when no solve ran) */
    int32_t polyfit_info;        /* out: SHOPT_POLY_* bits of the last round                       */
    int32_t stamps_f32;          /* in, shopt_stage_run / _comm only: `stamps` is the raw FITS VIGNET cube in float32
                                    (r, c, n_local); it is uploaded as such (half the bytes) and ingested on the device:
                                    nanMask (masks.jl:5-17) and, with unity_sum, ./ sum(nanToZero(.))                  */
    const int64_t *perm_dev;     /* in, optional: DEVICE copy of perm (saves the upload per call); perm may then be NULL */
    int32_t *model_status, *model_iters;       /* out, optional (n_local): SHOPT_FIT_* and iterations of the model fit   */
    int32_t *learned_status, *learned_iters;   /* out, optional (n_local): ... of the learned fit                        */
} shopt_stage_io;

/* Host logic of the split, exposed for sharded hosts and tests (shopt.jl:472-478): flags (n_global) = SHOPT_FLAG_* of
 * every star after the gather; train_out receives the first round(training_ratio * n_survivors) (half to even, like
 * Julia's round(Int, .)) survivors (flags & 7 == 0) in `perm` order; returns their count, or a negative SHOPT_E_*. */
int64_t shopt_training_split(const int32_t *flags, int64_t n_global, const int64_t *perm, double training_ratio,
                             int64_t *train_out);

/* HOST restatement of the sharded stage's exchange protocol (the device path runs the same steps as kernels,
 * csrc/select.cu): survivor bits along the permutation + 3 counters in ONE sum-all-reduce, then the training list of
 * this shard; worst_10_percent (shopt.jl:513-520) as an 8-pass radix select whose 4096-bin histogram is the only thing
 * ranks exchange.  Usable without a device: the world_size-2 gloo tests drive them.  allreduce: in-place sum over the
 * ranks of a HOST buffer of `count` elements, dtype 0 = float64, 1 = int32; NULL = single rank. */
typedef int (*shopt_host_allreduce_fn)(void *buf, int64_t count, int dtype, void *user);
int64_t shopt_training_split_sharded_host(const int32_t *flags_local, int64_t n_local, int64_t offset, int64_t n_global,
                                          const int64_t *perm, double training_ratio, shopt_host_allreduce_fn allreduce,
                                          void *user, int64_t *train_idx_out, uint32_t *train_pos_out, int64_t *counts_out);
int64_t shopt_select_worst_sharded_host(const double *err_local, const uint32_t *pos_local, int64_t n_local, int64_t m,
                                        shopt_host_allreduce_fn allreduce, void *user, uint8_t *reject_out);
/* per-pixel median(filter(!isnan, z)) (interpolate.jl:119,128,140) over the training stars of EVERY rank: z_local
 * (n_local, npx) row-major; med_out (npx), identical on every rank.  Radix select with a 256-bin histogram exchange per
 * pass, the protocol of the sharded stage's median_constraint / NaN -> median repair (csrc/polyfit.cu shard_med_*). */
int shopt_median_sharded_host(const double *z_local, int64_t n_local, int64_t npx, shopt_host_allreduce_fn allreduce,
                              void *user, double *med_out);

/* device-resident shard; comm == NULL runs a single GPU.  Blocks until the stage is done; the control flow between
 * the kernels (survivor counts, training split, worst_10_percent, list compaction) runs on the device, the call
 * synchronises the stream twice and allocates nothing after the first call of a given size.
 * Sharded runs serve the same options as one GPU: median_constraint and the NaN -> median repair (interpolate.jl:118-130)
 * take the per-pixel median over the training stars of EVERY shard by a radix select whose 256-bin histograms are the
 * only thing ranks exchange (16 int32 sum-all-reduces per 16 384-pixel tile), and cost one extra stream synchronisation
 * per round (every rank has to learn whether any pixel's all-reduced V'P is NaN). */
int shopt_stage_run_dev(const shopt_params *p, shopt_stage_io *io, shopt_allreduce_fn comm, void *comm_user, void *stream);
/* host buffers, one GPU: H2D (chunked on a copy stream, overlapped with the per-star kernels), stage, D2H */
int shopt_stage_run(const shopt_params *p, shopt_stage_io *io);
/* host buffers of ONE SHARD of a sharded run (one process per GPU): as shopt_stage_run, with the all-reduce callback of
 * shopt_stage_run_dev (enqueued on the library's stream, which is what `stream` of the callback carries). */
int shopt_stage_run_comm(const shopt_params *p, shopt_stage_io *io, shopt_allreduce_fn comm, void *comm_user);

#ifdef __cplusplus
}
#endif
#endif

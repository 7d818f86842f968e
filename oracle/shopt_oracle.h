/*
 * shopt_oracle.h -- CPU restatement (FP64, plain C) of ShOpt.jl's per-star PSF fitting stage.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker / the timed CPU baseline.  The product path (shopt_b200/) never
 * links, imports or executes anything from here and has no CPU fallback.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path
 * (SURVEY.md section 4 / 8c), Julia is not installed in the build container, and the numerics
 * live in un-vendored Julia packages pinned by shopt_env/Manifest.toml (Optim 1.7.8,
 * LineSearches 7.2.0, NLSolversBase 7.8.3, ForwardDiff 0.10.36, MultivariateStats 0.10.2,
 * LinearAlgebra/OpenBLAS).  This file restates their published algorithms (SURVEY.md
 * Appendix A) and anchors on the reference's own call sites; it is cross-checked in tests/
 * against independent numpy/scipy implementations and analytic known-answer cases.
 *
 * Array conventions follow Julia: column-major, stamp pixel (u,v) 1-based lives at
 * offset (u-1) + (v-1)*r.  Batches are (r, c, N) column-major, i.e. star i at stamps + i*r*c.
 * All indices crossing this C interface are 0-based.
 */
#ifndef SHOPT_ORACLE_H
#define SHOPT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes of one analytic fit (shared meaning with include/shopt_b200.h) */
#define ORC_FIT_CONVERGED      0  /* x-, f- or g-convergence (Optim assess_convergence)            */
#define ORC_FIT_MAXITER        1  /* 1000 iterations reached; minimizer = last iterate              */
#define ORC_FIT_LS_FAILED      2  /* LineSearchException: step alpha of the exception taken, stop   */
#define ORC_FIT_NONFINITE      3  /* non-finite value/gradient: stop, keep the current x            */
#define ORC_FIT_THREW         -1  /* an @assert/error() inside LineSearches fired -> shopt.jl catch */

/* ---- analytic profile: radialProfiles.jl:5-9, analyticLBFGS.jl:8-101 ---- */
double orc_fgaussian(double u, double v, double g1, double g2, double s, double uc, double vc);
/* radialProfiles.jl:11-24 with the quadgk integral in closed form */
double orc_fkolmogorov(double u, double v, double g1, double g2, double s, double uc, double vc);
/* radial profile used by orc_cost / orc_grad / orc_lbfgs_fit: 0 fGaussian (default; the reference's), 1 fkolmogorov */
void orc_set_profile(int profile);
int orc_get_profile(void);
double orc_cost(const double *stamp, int r, int c, int n_a, const double x[3]);
void   orc_grad(const double *stamp, int r, int c, int n_a, const double x[3], double g[3]);
void   orc_e_to_g(double e1, double e2, double *g1, double *g2);

/* Optim.optimize(cost, g!, x0, LBFGS(), Options(g_tol)) for one star (shopt.jl:139-153).
 * out_x = minimizer [sigma,e1,e2]; out_p = [s,g1,g2] after the shopt.jl post-processing
 * (zeros when status == ORC_FIT_THREW).  counts[0]=iterations, counts[1]=phi-dphi evaluations
 * actually executed (cache misses), counts[2]=line-search trial points. */
int orc_lbfgs_fit(const double *stamp, int r, int c, int n_a, const double x0[3],
                  double g_tol, int max_iter, double out_x[3], double out_p[3], int counts[3]);

/* loops shopt.jl:130-164 / 357-397 over a (r,c,N) batch; nthreads<=1 -> serial like the reference */
void orc_fit_analytic_batch(const double *stamps, int r, int c, int64_t N, int n_a,
                            const double *x0 /*3 x N*/, double g_tol, int max_iter,
                            double *s, double *g1, double *g2, int32_t *status, int32_t *iters,
                            int nthreads);

/* s-blacklist / failed-star selection shopt.jl:166-171, 384-386: returns count, idx ascending */
int64_t orc_s_out_of_bounds(const double *s, int64_t N, double s_lo, double s_hi, int64_t *idx);

/* ---- pixel-grid fit: masks.jl:38-50, pca.jl:1-18, dataPreprocessing.jl:94-139 ---- */
void orc_nan_to_zero(const double *in, int64_t n, double *out);
int  orc_pca_image(const double *X, int r, int c, int ncomp, double pratio, double *out, int *k_out);
void orc_lanczos_kernel2d(int a, double *kernel /*(2a+1)^2 col-major*/);
void orc_smooth(const double *in, int r, int c, int a, double *out);
/* PCA-mode loop shopt.jl:281-314 (including the unity_sum 1-argument smooth() quirk) */
void orc_pixel_grid_pca_batch(const double *stamps, int r, int c, int64_t N, int ncomp, double pratio,
                              int lanczos_a, int unity_sum, double *pgf, int32_t *k_out,
                              int32_t *status, int nthreads);

/* ---- interpolation: interpolate.jl:89-147, shopt.jl:472-577 ---- */
int    orc_num_terms(int degree);
void   orc_poly_terms(int degree, int *p, int *q);   /* polynomial_optimizer order (p outer, q inner) */
double orc_objective_function(const double *coef, double x, double y, int degree);
/* A \ z by column-pivoted Householder QR (+ min-norm when rank deficient). z is modified in
 * place when NaNs are replaced, like the reference (interpolate.jl:126-130). */
void   orc_polynomial_optimizer(int degree, const double *x, const double *y, double *z, int64_t n,
                                int median_constraint, double *coef);
double orc_median(const double *z, int64_t n);       /* of the non-NaN entries, Statistics.median */
int64_t orc_worst_10_percent(const double *err, int64_t n, int sigma_clip, int64_t *idx_out);
/* full driver of shopt.jl:534-577.  train_idx: 0-based indices into the N stars (the randperm
 * split is an INPUT).  poly_out: (r,c,K) column-major.  mse_out: last round's training errors in
 * that round's training order; kept_idx/n_kept: training set after the last applied rejection
 * (original star indices); n_per_round[t]: training-set size used in round t. */
void orc_polyfit(const double *pgf, int r, int c, int64_t N, const double *u, const double *v,
                 const int64_t *train_idx, int64_t n_train, int degree, int iterations,
                 int median_constraint, int sigma_clip, double *poly_out, double *mse_out,
                 int64_t *kept_idx, int64_t *n_kept, int64_t *n_per_round, int nthreads);
/* reader.jl:40-68  p(u,v,polMatrix,degree): evaluate + unit-sum normalise */
void orc_psf_at(const double *poly, int r, int c, int degree, double u, double v, double *star);

int orc_max_threads(void);
/* 1: every value+gradient evaluation also runs the separate Float64 cost() pass, as NLSolversBase's
 * fg!(G,x) = (g!(G,x); f(x)) does in the reference (bit-identical value; only the CPU time differs). */
void orc_set_reference_shaped(int on);

#ifdef __cplusplus
}
#endif
#endif

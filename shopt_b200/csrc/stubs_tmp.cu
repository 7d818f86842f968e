// temporary: entry points not implemented yet return SHOPT_E_UNSUPPORTED (removed as kernels land)
#include "common.cuh"
extern "C" {
#ifndef HAVE_PCA
int shopt_pca_batch(const double *, int, int, int64_t, int, double, int, int, double *, int32_t *, int32_t *) { return SHOPT_E_UNSUPPORTED; }
int shopt_pca_batch_dev(const double *, int, int, int64_t, int, double, int, int, double *, int32_t *, int32_t *, void *) { return SHOPT_E_UNSUPPORTED; }
int shopt_smooth_batch(const double *, int, int, int64_t, int, int, double *) { return SHOPT_E_UNSUPPORTED; }
int shopt_smooth_batch_dev(const double *, int, int, int64_t, int, int, double *, void *) { return SHOPT_E_UNSUPPORTED; }
#endif
#ifndef HAVE_POLY
int shopt_polyfit(const double *, int, int, int64_t, const double *, const double *, const int64_t *, int64_t, int, int, int, int, double *, double *, int64_t *, int64_t *, int64_t *) { return SHOPT_E_UNSUPPORTED; }
int64_t shopt_polyfit_acc_len(int r, int c, int degree) { int K = (degree + 1) * (degree + 2) / 2; return (int64_t)K * K + (int64_t)K * r * c; }
int shopt_polyfit_normal_dev(const double *, int, int, const double *, const double *, const int64_t *, int64_t, int, double *, void *) { return SHOPT_E_UNSUPPORTED; }
int shopt_polyfit_solve_dev(const double *, int, int, int, double *, int32_t *, void *) { return SHOPT_E_UNSUPPORTED; }
int shopt_polyfit_mse_dev(const double *, int, int, const double *, const double *, const int64_t *, int64_t, int, const double *, double *, void *) { return SHOPT_E_UNSUPPORTED; }
int64_t shopt_worst_10_percent(const double *, int64_t, int, int64_t *) { return SHOPT_E_UNSUPPORTED; }
int shopt_psf_at_batch(const double *, int, int, int, const double *, const double *, int64_t, double *) { return SHOPT_E_UNSUPPORTED; }
int shopt_psf_at_batch_dev(const double *, int, int, int, const double *, const double *, int64_t, double *, void *) { return SHOPT_E_UNSUPPORTED; }
#endif
}

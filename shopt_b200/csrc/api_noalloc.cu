// api_noalloc.cu -- shopt_polyfit_mse_noalloc_dev: the per-star reconstruction error of the interpolation step
// (mse_kernel + mse_reduce_kernel of polyfit.cu) for callers that cannot allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the
// reference.
//
// The per-tile partial sums -- the scratch the ordinary entry point (shopt_polyfit_mse_dev) takes from the context's
// grow-only block -- live in a file-scope __device__ array of SHOPT_NOALLOC_MSE_SCRATCH_BYTES.  Every GPU that loads the
// library holds that array until the process exits.  The entry point creates no context (get_ctx allocates on first
// use), calls no cudaMalloc / cudaFree / allocating library routine, and rejects problems whose partial sums do not fit.
//
// All calls share the one array.  It keeps the previous call's contents, which is harmless here: mse_kernel WRITES every
// partial sum part[tile][t], t < n, that mse_reduce_kernel then reads, so nothing has to start at zero and nothing is
// cleared.  Two calls that overlap on the GPU (different streams, or a graph replayed concurrently with itself) race on
// the array: the caller serialises them.
#include <cstdio>
#include <mutex>

#include "api_ctx.h"

namespace shopt {

constexpr size_t NOALLOC_MSE_DOUBLES = (size_t)SHOPT_NOALLOC_MSE_SCRATCH_BYTES / sizeof(double);
static __device__ __align__(256) double g_noalloc_mse_part[NOALLOC_MSE_DOUBLES];

// In host code g_noalloc_mse_part is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static std::mutex g_noalloc_mu;
static double *g_noalloc_ptr[64] = {};

static int noalloc_scratch(double **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_last_error("cudaGetDevice", e); (void)cudaGetLastError(); return SHOPT_E_CUDA; }
    if (dev < 0 || dev >= 64) { set_last_error_text("shopt_polyfit_mse_noalloc_dev: device ordinal out of range"); return SHOPT_E_BADARG; }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_mse_part);
        if (e != cudaSuccess || !p) {
            set_last_error("cudaGetSymbolAddress(g_noalloc_mse_part)", e);
            (void)cudaGetLastError();
            return SHOPT_E_CUDA;
        }
        g_noalloc_ptr[dev] = static_cast<double *>(p);
    }
    *out = g_noalloc_ptr[dev];
    return SHOPT_OK;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int64_t shopt_polyfit_mse_noalloc_max_stars(int r, int c, int degree) {
    if (r < 1 || c < 1 || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    const size_t per_star = polyfit_mse_part_len((long long)r * c, 1, degree);   // one partial sum per pixel tile
    return (int64_t)(NOALLOC_MSE_DOUBLES / per_star);
}

int shopt_polyfit_mse_noalloc_dev(const double *d_pgf, int r, int c, const double *d_u, const double *d_v,
                                  const int64_t *d_idx, int64_t n, int degree, const double *d_poly, double *d_mse,
                                  void *stream) {
    if (!d_pgf || !d_u || !d_v || !d_poly || !d_mse || r < 1 || c < 1 || n < 0 || degree < 0 || degree > 8)
        return SHOPT_E_BADARG;
    const int64_t max_stars = shopt_polyfit_mse_noalloc_max_stars(r, c, degree);
    if (n > max_stars) {
        char text[256];
        snprintf(text, sizeof text,
                 "shopt_polyfit_mse_noalloc_dev: %lld stars of %d x %d pixels at degree %d exceed the static scratch of %lld "
                 "bytes (at most %lld stars of this shape per call); use shopt_polyfit_mse_dev or split the list",
                 (long long)n, r, c, degree, (long long)SHOPT_NOALLOC_MSE_SCRATCH_BYTES, (long long)max_stars);
        set_last_error_text(text);
        return SHOPT_E_UNSUPPORTED;
    }
    if (n == 0) return SHOPT_OK;
    double *part = nullptr;
    TRY(noalloc_scratch(&part));
    StarList sl; sl.idx = (const long long *)d_idx; sl.n = n;
    return polyfit_mse_launch(d_pgf, (long long)r * c, d_u, d_v, sl, degree, d_poly, d_mse, part,
                              (size_t)SHOPT_NOALLOC_MSE_SCRATCH_BYTES, (cudaStream_t)stream);
}

}  // extern "C"

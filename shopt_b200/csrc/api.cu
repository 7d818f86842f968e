// api.cu -- the C ABI of libshopt_b200.so (include/shopt_b200.h): argument checks, device buffers,
// H2D/D2H staging and stream handling around the kernels in lbfgs_fit.cu / pca.cu / polyfit.cu.
// No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <vector>

#include "api_ctx.h"

namespace shopt {

static thread_local char g_err[512] = "";
void set_last_error_text(const char *text) { snprintf(g_err, sizeof g_err, "%s", text); }
void set_last_error(const char *what, cudaError_t e) {
    snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
}
int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

int Ctx::need_scratch(size_t bytes) {
    if (bytes <= scratch_bytes) return SHOPT_OK;
    if (scratch) { cudaDeviceSynchronize(); cudaFree(scratch); scratch = nullptr; scratch_bytes = 0; }
    cudaError_t e = cudaMalloc(&scratch, bytes);
    if (e != cudaSuccess) { set_last_error("cudaMalloc(scratch)", e); (void)cudaGetLastError(); return SHOPT_E_NOMEM; }
    scratch_bytes = bytes;
    return SHOPT_OK;
}
int Ctx::need_cache(int i, size_t bytes, void **out) {
    if (i < 0 || i >= NCACHE) return SHOPT_E_BADARG;
    if (bytes == 0) bytes = 16;
    if (bytes > cache_bytes[i]) {
        if (cache[i]) { cudaDeviceSynchronize(); cudaFree(cache[i]); cache[i] = nullptr; cache_bytes[i] = 0; }
        cudaError_t e = cudaMalloc(&cache[i], bytes);
        if (e != cudaSuccess) { cache[i] = nullptr; set_last_error("cudaMalloc(cache)", e); (void)cudaGetLastError(); return SHOPT_E_NOMEM; }
        cache_bytes[i] = bytes;
    }
    *out = cache[i];
    return SHOPT_OK;
}
void Ctx::release_cache() {
    cudaDeviceSynchronize();
    for (int i = 0; i < NCACHE; i++) { if (cache[i]) cudaFree(cache[i]); cache[i] = nullptr; cache_bytes[i] = 0; }
    if (scratch) { cudaFree(scratch); scratch = nullptr; scratch_bytes = 0; }
}
static thread_local Ctx g_ctx[16];

int get_ctx(Ctx **out) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1) {
        snprintf(g_err, sizeof g_err, "no CUDA device");
        (void)cudaGetLastError();
        return SHOPT_E_NODEVICE;
    }
    int dev = 0;
    SHOPT_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 16) return SHOPT_E_UNSUPPORTED;
    Ctx &c = g_ctx[dev];
    if (c.dev != dev) {
        SHOPT_CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
        SHOPT_CUDA_TRY(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 4; i++) SHOPT_CUDA_TRY(cudaEventCreate(&c.ev[i]));
        for (int i = 0; i < Ctx::NCHUNK_EV; i++) SHOPT_CUDA_TRY(cudaEventCreateWithFlags(&c.chunk_ev[i], cudaEventDisableTiming));
        SHOPT_CUDA_TRY(cudaMalloc(&c.counter, 64));
        c.dev = dev;
    }
    *out = &c;
    return SHOPT_OK;
}

int finish_timing(Ctx *c) {
    SHOPT_CUDA_TRY(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 3; i++) {
        float ms = 0.f;
        SHOPT_CUDA_TRY(cudaEventElapsedTime(&ms, c->ev[i], c->ev[i + 1]));
        c->ms[i] = ms;
    }
    return SHOPT_OK;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int shopt_b200_version(void) { return 100; }

int shopt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    return n;
}
int shopt_set_device(int dev) {
    if (shopt_device_count() < 1) return SHOPT_E_NODEVICE;
    SHOPT_CUDA_TRY(cudaSetDevice(dev));
    return SHOPT_OK;
}
const char *shopt_last_error(void) { return g_err; }
int shopt_last_timing(double ms[3]) {
    if (!ms) return SHOPT_E_BADARG;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) { (void)cudaGetLastError(); return SHOPT_E_NODEVICE; }
    for (int i = 0; i < 3; i++) ms[i] = g_ctx[dev].ms[i];
    return SHOPT_OK;
}

int shopt_release_buffers(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) { (void)cudaGetLastError(); return SHOPT_E_NODEVICE; }
    if (g_ctx[dev].dev == dev) g_ctx[dev].release_cache();
    return SHOPT_OK;
}

void shopt_default_params(shopt_params *p) {
    if (!p) return;
    p->n_a = 64; p->g_tol_model = 1e-6; p->g_tol_learned = 1e-6; p->max_iter = 1000;
    p->s_lo = 0.075; p->s_hi = 1.0; p->pca_terms = 50; p->pratio = 0.99; p->lanczos = 3; p->unity_sum = 0;
    p->degree = 3; p->iterations = 3; p->training_ratio = 0.8; p->median_constraint = 0; p->sigma_clip = 0;
    p->mode = SHOPT_MODE_PCA; p->chisq_fill = 0.0;
}

// ---- (1) analytic fit ---------------------------------------------------------------------
int shopt_fit_analytic_batch_dev(const double *d_stamps, int r, int c, int64_t N, int n_a, const double *d_x0,
                                 double g_tol, int max_iter, double *d_s, double *d_g1, double *d_g2,
                                 int32_t *d_status, int32_t *d_iters, void *stream) {
    Ctx *ctx; TRY(get_ctx(&ctx));
    return fit_analytic_launch(d_stamps, r, c, (long long)N, n_a, d_x0, g_tol, max_iter, d_s, d_g1, d_g2, d_status,
                               d_iters, ctx->next_counter(), (cudaStream_t)stream);
}

int shopt_fit_analytic_batch(const double *stamps, int r, int c, int64_t N, int n_a, const double *x0, double g_tol,
                             int max_iter, double *s, double *g1, double *g2, int32_t *status, int32_t *iters) {
    if (!stamps || !x0 || !s || !g1 || !g2) return SHOPT_E_BADARG;
    if (r < 1 || c < 1 || n_a < 1 || n_a > r || n_a > c || N < 0 || max_iter < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    const size_t rc = (size_t)r * c, n = (size_t)N;
    DevBuf dst, dx0, dout, dint;
    TRY(dst.alloc(n * rc * 8)); TRY(dx0.alloc(n * 24)); TRY(dout.alloc(n * 24)); TRY(dint.alloc(n * 8));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(dst.p, stamps, n * rc * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(dx0.p, x0, n * 24, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    double *ds = dout.as<double>();
    int32_t *di = dint.as<int32_t>();
    TRY(fit_analytic_launch(dst.as<double>(), r, c, (long long)N, n_a, dx0.as<double>(), g_tol, max_iter, ds, ds + n,
                            ds + 2 * n, di, di + n, ctx->next_counter(), st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(s, ds, n * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(g1, ds + n, n * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(g2, ds + 2 * n, n * 8, cudaMemcpyDeviceToHost, st));
    if (status) SHOPT_CUDA_TRY(cudaMemcpyAsync(status, di, n * 4, cudaMemcpyDeviceToHost, st));
    if (iters) SHOPT_CUDA_TRY(cudaMemcpyAsync(iters, di + n, n * 4, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

int shopt_analytic_cost_grad_batch_dev(const double *d_stamps, int r, int c, int64_t N, int n_a, const double *d_x,
                                       double *d_f, double *d_g, void *stream) {
    if (!d_f || !d_g) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    return fit_analytic_launch(d_stamps, r, c, (long long)N, n_a, d_x, 0.0, 0, nullptr, nullptr, nullptr, nullptr,
                               nullptr, ctx->next_counter(), (cudaStream_t)stream, d_f, d_g);
}

int shopt_analytic_cost_grad_batch(const double *stamps, int r, int c, int64_t N, int n_a, const double *x, double *f,
                                   double *g) {
    if (!stamps || !x || !f || !g) return SHOPT_E_BADARG;
    if (r < 1 || c < 1 || n_a < 1 || n_a > r || n_a > c || N < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    const size_t rc = (size_t)r * c, n = (size_t)N;
    DevBuf dst, dx, dout;
    TRY(dst.alloc(n * rc * 8)); TRY(dx.alloc(n * 24)); TRY(dout.alloc(n * 32));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaMemcpyAsync(dst.p, stamps, n * rc * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(dx.p, x, n * 24, cudaMemcpyHostToDevice, st));
    double *df = dout.as<double>();
    TRY(fit_analytic_launch(dst.as<double>(), r, c, (long long)N, n_a, dx.as<double>(), 0.0, 0, nullptr, nullptr,
                            nullptr, nullptr, nullptr, ctx->next_counter(), st, df, df + n));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(f, df, n * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(g, df + n, n * 24, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    return SHOPT_OK;
}

int64_t shopt_s_out_of_bounds(const double *s, int64_t N, double s_lo, double s_hi, int64_t *idx) {
    if (!s || !idx || N < 0) return SHOPT_E_BADARG;
    int64_t n = 0;
    for (int64_t i = 0; i < N; i++)
        if (s[i] < s_lo || s[i] > s_hi) idx[n++] = i;   // shopt.jl:167-171
    return n;
}

}  // extern "C"

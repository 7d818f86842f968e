// api.cu -- the C ABI of libshopt_b200.so (include/shopt_b200.h): argument checks, device buffers,
// H2D/D2H staging and stream handling around the kernels in lbfgs_fit.cu / pca.cu / polyfit.cu.
// No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <numeric>
#include <vector>

#include "api_ctx.h"

namespace shopt {

static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
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
// grow-only block: freed (after the device has drained) only when a larger one is needed
static int grow_block(void **p, size_t *have, size_t bytes, const char *what) {
    if (bytes <= *have) return SHOPT_OK;
    if (*p) { cudaDeviceSynchronize(); cudaFree(*p); *p = nullptr; *have = 0; }
    bytes = (bytes + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) { *p = nullptr; set_last_error(what, e); (void)cudaGetLastError(); return SHOPT_E_NOMEM; }
    *have = bytes;
    return SHOPT_OK;
}
int Ctx::need_arena(size_t bytes) { return grow_block(&arena, &arena_bytes, bytes, "cudaMalloc(arena)"); }
int Ctx::need_arena2(size_t bytes) { return grow_block(&arena2, &arena2_bytes, bytes, "cudaMalloc(arena2)"); }
int Ctx::need_pinned(size_t bytes) {
    if (bytes <= pinned_bytes) return SHOPT_OK;
    if (pinned) { cudaDeviceSynchronize(); cudaFreeHost(pinned); pinned = nullptr; pinned_bytes = 0; }
    bytes = (bytes + 65535) & ~(size_t)65535;
    cudaError_t e = cudaHostAlloc(&pinned, bytes, cudaHostAllocDefault);
    if (e != cudaSuccess) { pinned = nullptr; set_last_error("cudaHostAlloc", e); (void)cudaGetLastError(); return SHOPT_E_NOMEM; }
    pinned_bytes = bytes;
    return SHOPT_OK;
}
void Ctx::release_cache() {
    cudaDeviceSynchronize();
    for (int i = 0; i < NCACHE; i++) { if (cache[i]) cudaFree(cache[i]); cache[i] = nullptr; cache_bytes[i] = 0; }
    if (scratch) { cudaFree(scratch); scratch = nullptr; scratch_bytes = 0; }
    if (arena) { cudaFree(arena); arena = nullptr; arena_bytes = 0; }
    if (arena2) { cudaFree(arena2); arena2 = nullptr; arena2_bytes = 0; }
    if (pinned) { cudaFreeHost(pinned); pinned = nullptr; pinned_bytes = 0; }
}
// One context per device for the whole process (not per thread: a second calling thread would silently get its own
// streams and its own tens of GB of cached blocks).  Calls on ONE device must not overlap; the reference calls the
// stage from its main thread only.
static Ctx g_ctx[16];
static std::mutex g_ctx_mutex;

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
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
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

// FP64 FMA peak: 16 independent chains per thread, 8 CTAs of 256 threads per SM
__global__ void __launch_bounds__(256) dfma_kernel(double *out, int iters, double a, double b) {
    double x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = (double)threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = fma(x[i], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += x[i];
    if (s == 12345.678) out[0] = s;      // never true: keeps the chains alive
}

}  // namespace shopt

using namespace shopt;

extern "C" {

// Measured FP64 FMA throughput of the current device (TFLOP/s), the compute roofline of fit_kernel / pca_kernel
// (tcgen05 has no FP64 path; MEASURED_PEAKS.json holds no FP64 figure).
int shopt_bench_dfma(double *tflops_out) {
    if (!tflops_out) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    TRY(ctx->need_arena2(256));
    const int iters = 20000, grid = sm_count() * 8;
    cudaStream_t st = ctx->stream;
    dfma_kernel<<<grid, 256, 0, st>>>((double *)ctx->arena2, 1000, 0.999999, 1e-9);     // warm-up
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    dfma_kernel<<<grid, 256, 0, st>>>((double *)ctx->arena2, iters, 0.999999, 1e-9);
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    float ms = 0.f;
    SHOPT_CUDA_TRY(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    *tflops_out = 2.0 * 16.0 * iters * 256.0 * grid / (ms * 1e-3) / 1e12;
    return SHOPT_OK;
}

int shopt_b200_version(void) { return 200; }
int64_t shopt_launch_count(void) { return (int64_t)g_launches.load(); }

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
    p->mode = SHOPT_MODE_PCA; p->chisq_fill = 0.0; p->profile = SHOPT_PROFILE_GAUSSIAN;
}

// ---- (1) analytic fit ---------------------------------------------------------------------
int shopt_fit_profile_batch_dev(int profile, const double *d_stamps, int r, int c, int64_t N, int n_a, const double *d_x0,
                                double g_tol, int max_iter, double *d_s, double *d_g1, double *d_g2,
                                int32_t *d_status, int32_t *d_iters, void *stream) {
    Ctx *ctx; TRY(get_ctx(&ctx));
    return fit_analytic_launch(d_stamps, r, c, (long long)N, n_a, d_x0, g_tol, max_iter, d_s, d_g1, d_g2, d_status,
                               d_iters, ctx->next_counter(), (cudaStream_t)stream, nullptr, nullptr, profile);
}
int shopt_fit_analytic_batch_dev(const double *d_stamps, int r, int c, int64_t N, int n_a, const double *d_x0,
                                 double g_tol, int max_iter, double *d_s, double *d_g1, double *d_g2,
                                 int32_t *d_status, int32_t *d_iters, void *stream) {
    return shopt_fit_profile_batch_dev(SHOPT_PROFILE_GAUSSIAN, d_stamps, r, c, N, n_a, d_x0, g_tol, max_iter, d_s, d_g1, d_g2,
                                       d_status, d_iters, stream);
}

int shopt_fit_analytic_batch(const double *stamps, int r, int c, int64_t N, int n_a, const double *x0, double g_tol,
                             int max_iter, double *s, double *g1, double *g2, int32_t *status, int32_t *iters) {
    return shopt_fit_profile_batch(SHOPT_PROFILE_GAUSSIAN, stamps, r, c, N, n_a, x0, g_tol, max_iter, s, g1, g2, status, iters);
}
int shopt_fit_profile_batch(int profile, const double *stamps, int r, int c, int64_t N, int n_a, const double *x0, double g_tol,
                            int max_iter, double *s, double *g1, double *g2, int32_t *status, int32_t *iters) {
    if (profile < 0 || profile > 1) return SHOPT_E_BADARG;
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
                            ds + 2 * n, di, di + n, ctx->next_counter(), st, nullptr, nullptr, profile));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(s, ds, n * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(g1, ds + n, n * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(g2, ds + 2 * n, n * 8, cudaMemcpyDeviceToHost, st));
    if (status) SHOPT_CUDA_TRY(cudaMemcpyAsync(status, di, n * 4, cudaMemcpyDeviceToHost, st));
    if (iters) SHOPT_CUDA_TRY(cudaMemcpyAsync(iters, di + n, n * 4, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

int shopt_profile_cost_grad_batch_dev(int profile, const double *d_stamps, int r, int c, int64_t N, int n_a,
                                      const double *d_x, double *d_f, double *d_g, void *stream) {
    if (!d_f || !d_g) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    return fit_analytic_launch(d_stamps, r, c, (long long)N, n_a, d_x, 0.0, 0, nullptr, nullptr, nullptr, nullptr,
                               nullptr, ctx->next_counter(), (cudaStream_t)stream, d_f, d_g, profile);
}
int shopt_analytic_cost_grad_batch_dev(const double *d_stamps, int r, int c, int64_t N, int n_a, const double *d_x,
                                       double *d_f, double *d_g, void *stream) {
    return shopt_profile_cost_grad_batch_dev(SHOPT_PROFILE_GAUSSIAN, d_stamps, r, c, N, n_a, d_x, d_f, d_g, stream);
}

int shopt_analytic_cost_grad_batch(const double *stamps, int r, int c, int64_t N, int n_a, const double *x, double *f,
                                   double *g) {
    return shopt_profile_cost_grad_batch(SHOPT_PROFILE_GAUSSIAN, stamps, r, c, N, n_a, x, f, g);
}
int shopt_profile_cost_grad_batch(int profile, const double *stamps, int r, int c, int64_t N, int n_a, const double *x,
                                  double *f, double *g) {
    if (profile < 0 || profile > 1) return SHOPT_E_BADARG;
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
                            nullptr, nullptr, nullptr, ctx->next_counter(), st, df, df + n, profile));
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

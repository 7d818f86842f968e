// api_pixelgrid.cu -- C ABI of the pixel-grid fit (include/shopt_b200.h section 2): PCA mode and smooth().
#include "api_ctx.h"

using namespace shopt;

// one buffer in, one buffer out, host side: H2D, launch, D2H with the context's timing events
template <class F>
static int host_map(const double *in, const double *in2, const double *in3, double *out, size_t bytes, F launch) {
    Ctx *ctx; TRY(get_ctx(&ctx));
    DevBuf din, din2, din3, dout;
    TRY(din.alloc(bytes)); TRY(dout.alloc(bytes));
    if (in2) TRY(din2.alloc(bytes));
    if (in3) TRY(din3.alloc(bytes));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(din.p, in, bytes, cudaMemcpyHostToDevice, st));
    if (in2) SHOPT_CUDA_TRY(cudaMemcpyAsync(din2.p, in2, bytes, cudaMemcpyHostToDevice, st));
    if (in3) SHOPT_CUDA_TRY(cudaMemcpyAsync(din3.p, in3, bytes, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    TRY(launch(din.as<double>(), in2 ? din2.as<double>() : nullptr, in3 ? din3.as<double>() : nullptr, dout.as<double>(), st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(out, dout.p, bytes, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

extern "C" {

int shopt_pca_batch_dev(const double *d_stamps, int r, int c, int64_t N, int ncomp, double pratio, int lanczos_a,
                        int unity_sum, double *d_pgf, int32_t *d_k_out, int32_t *d_status, void *stream) {
    if (!d_stamps || !d_pgf || r < 1 || c < 2 || N < 0 || ncomp < 1) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    TRY(ctx->need_scratch(pca_scratch_bytes(r, c, ncomp, (long long)N)));
    return pca_launch(d_stamps, r, c, (long long)N, ncomp, pratio, lanczos_a, unity_sum, d_pgf, d_k_out, d_status,
                      ctx->next_counter(), (double *)ctx->scratch, ctx->scratch_bytes, (cudaStream_t)stream);
}

int shopt_pca_batch(const double *stamps, int r, int c, int64_t N, int ncomp, double pratio, int lanczos_a, int unity_sum,
                    double *pgf, int32_t *k_out, int32_t *status) {
    if (!stamps || !pgf || r < 1 || c < 2 || N < 0 || ncomp < 1) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    const size_t rc = (size_t)r * c, n = (size_t)N;
    DevBuf din, dout, dint;
    TRY(din.alloc(n * rc * 8)); TRY(dout.alloc(n * rc * 8)); TRY(dint.alloc(n * 8));
    TRY(ctx->need_scratch(pca_scratch_bytes(r, c, ncomp, (long long)N)));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(din.p, stamps, n * rc * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    int32_t *di = dint.as<int32_t>();
    TRY(pca_launch(din.as<double>(), r, c, (long long)N, ncomp, pratio, lanczos_a, unity_sum, dout.as<double>(), di, di + n,
                   ctx->next_counter(), (double *)ctx->scratch, ctx->scratch_bytes, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(pgf, dout.p, n * rc * 8, cudaMemcpyDeviceToHost, st));
    if (k_out) SHOPT_CUDA_TRY(cudaMemcpyAsync(k_out, di, n * 4, cudaMemcpyDeviceToHost, st));
    if (status) SHOPT_CUDA_TRY(cudaMemcpyAsync(status, di + n, n * 4, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

int shopt_smooth_batch_dev(const double *d_stamps, int r, int c, int64_t N, int lanczos_a, int nan_to_zero, double *d_out,
                           void *stream) {
    Ctx *ctx; TRY(get_ctx(&ctx));
    return smooth_launch(d_stamps, r, c, (long long)N, lanczos_a, nan_to_zero, d_out, (cudaStream_t)stream);
}

int shopt_smooth_batch(const double *stamps, int r, int c, int64_t N, int lanczos_a, int nan_to_zero, double *out) {
    if (!stamps || !out || r < 1 || c < 1 || N < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    const size_t bytes = (size_t)N * r * c * 8;
    DevBuf din, dout;
    TRY(din.alloc(bytes)); TRY(dout.alloc(bytes));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(din.p, stamps, bytes, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    TRY(smooth_launch(din.as<double>(), r, c, (long long)N, lanczos_a, nan_to_zero, dout.as<double>(), st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(out, dout.p, bytes, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

int shopt_smooth_mode_batch_dev(const double *d_stamps, int r, int c, int64_t N, int lanczos_a, int unity_sum,
                                double *d_pgf, void *stream) {
    if (!d_stamps || !d_pgf || r < 1 || c < 1 || N < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (unity_sum)
        return elem_pgf_launch(d_stamps, nullptr, nullptr, 0.0, (long long)r * c, (long long)N, 0, 1, d_pgf, (cudaStream_t)stream);
    return smooth_launch(d_stamps, r, c, (long long)N, lanczos_a, 1, d_pgf, (cudaStream_t)stream);
}
int shopt_smooth_mode_batch(const double *stamps, int r, int c, int64_t N, int lanczos_a, int unity_sum, double *pgf) {
    if (!stamps || !pgf || r < 1 || c < 1 || N < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    return host_map(stamps, nullptr, nullptr, pgf, (size_t)N * r * c * 8,
                    [&](const double *a, const double *, const double *, double *o, cudaStream_t st) {
                        return shopt_smooth_mode_batch_dev(a, r, c, N, lanczos_a, unity_sum, o, st);
                    });
}

int shopt_chisq_batch_dev(const double *d_stamps, const double *d_weights, const double *d_x0, double fill, int r, int c,
                          int64_t N, int unity_sum, double *d_pgf, void *stream) {
    if (!d_stamps || !d_weights || !d_pgf || r < 1 || c < 1 || N < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    return elem_pgf_launch(d_stamps, d_weights, d_x0, fill, (long long)r * c, (long long)N, 1, unity_sum ? 1 : 0, d_pgf,
                           (cudaStream_t)stream);
}
int shopt_chisq_batch(const double *stamps, const double *weights, const double *x0, double fill, int r, int c, int64_t N,
                      int unity_sum, double *pgf) {
    if (!stamps || !weights || !pgf || r < 1 || c < 1 || N < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    return host_map(stamps, weights, x0, pgf, (size_t)N * r * c * 8,
                    [&](const double *a, const double *w, const double *z, double *o, cudaStream_t st) {
                        return shopt_chisq_batch_dev(a, w, z, fill, r, c, N, unity_sum, o, st);
                    });
}

}  // extern "C"

// api_pixelgrid.cu -- C ABI of the pixel-grid fit (include/shopt_b200.h section 2): PCA mode and smooth().
#include "api_ctx.h"

using namespace shopt;

extern "C" {

int shopt_pca_batch_dev(const double *d_stamps, int r, int c, int64_t N, int ncomp, double pratio, int lanczos_a,
                        int unity_sum, double *d_pgf, int32_t *d_k_out, int32_t *d_status, void *stream) {
    if (!d_stamps || !d_pgf || r < 1 || c < 2 || N < 0 || ncomp < 1) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    TRY(ctx->need_scratch(pca_scratch_bytes(r, c, ncomp)));
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
    TRY(ctx->need_scratch(pca_scratch_bytes(r, c, ncomp)));
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

}  // extern "C"

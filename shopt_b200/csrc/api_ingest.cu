// api_ingest.cu -- the per-star work of cataloging() that precedes the fitting stage.
//
// Replaces, in shopt/dataPreprocessing.jl:342-397 (FITS decoding itself stays with astropy / CFITSIO):
//   nanMask(vignet)                                   masks.jl:5-17        x < -1000 -> NaN (SExtractor's mask sentinel)
//   ./ sum(nanToZero(nanMask(vignet)))  if unity_sum  :344-346
//   outliers_filter(snr, ., snrCutoff)                outliers.jl:16-34    (host: shopt_snr_cut)
//   new_img_dim < size: get_middle_nxn crop           :207-216, :384       then nm(.) / sum(nz(nm(.))), smooth(., lanczos)
//   new_img_dim > size: bilinear oversample_image     :141-162, :391       then nm(.) / sum(nz(nm(.))), smooth(., lanczos)
// One CTA per star: HBM-bound, coalesced loads of the raw float32 / float64 vignet, one write of the float64 stamp.
#include <algorithm>
#include <cmath>
#include <vector>

#include "api_ctx.h"

namespace shopt {

struct IngestArgs {
    const void *in;
    double *out;
    long long N;
    int f32, r0, c0, nd, unity, resize;
};

__device__ __forceinline__ double ingest_load(const IngestArgs &a, long long star, int i, int j) {   // nanMask on the way in
    const long long p = star * (long long)a.r0 * a.c0 + i + (long long)j * a.r0;
    const double x = a.f32 ? (double)__ldg(reinterpret_cast<const float *>(a.in) + p)
                           : __ldg(reinterpret_cast<const double *>(a.in) + p);
    return (x < -1000.0) ? nan("") : x;
}
__device__ __forceinline__ double ntz(double x) { return isnan(x) ? 0.001 : x; }   // masks.jl:43
__device__ __forceinline__ double nm(double x) { return (x < -1000.0) ? nan("") : x; }

__global__ void __launch_bounds__(256) ingest_kernel(const IngestArgs a) {
    __shared__ double sred[2 * 8];
    int buf = 0;
    const int nd = a.nd;
    const long long npx = a.resize ? (long long)nd * nd : (long long)a.r0 * a.c0;
    for (long long star = blockIdx.x; star < a.N; star += gridDim.x) {
        double *out = a.out + star * npx;
        // the catalogue entry: nanMask(vignet) [./ sum(nanToZero(nanMask(vignet)))]
        double tot0 = 1.0;
        if (a.unity) {
            double part = 0.0;
            for (long long p = threadIdx.x; p < (long long)a.r0 * a.c0; p += 256)
                part += ntz(ingest_load(a, star, (int)(p % a.r0), (int)(p / a.r0)));
            double v1[1] = {part};
            cta_sum<1, 8>(v1, sred, buf);
            tot0 = v1[0];
        }
        if (!a.resize) {
            for (long long p = threadIdx.x; p < npx; p += 256) {
                const double x = ingest_load(a, star, (int)(p % a.r0), (int)(p / a.r0));
                out[p] = a.unity ? x / tot0 : x;
            }
            continue;
        }
        // crop or bilinear resample of the catalogue entry, then nm(.) / sum(nz(nm(.)))  (smoothing follows in smooth_kernel)
        double part = 0.0;
        const int rs = (a.r0 - nd) / 2, cs = (a.c0 - nd) / 2;                 // get_middle_nxn: div(rows - n, 2) + 1, 1-based
        const double scale = 1.0 / ((double)nd / (double)a.r0);               // oversample_image: 1/(new_dim / size(image)[1])
        for (long long p = threadIdx.x; p < npx; p += 256) {
            const int y = (int)(p % nd), x = (int)(p / nd);                   // out[y, x], y the fast (row) index
            double val;
            if (nd < a.r0) {
                val = ingest_load(a, star, rs + y, cs + x);
                if (a.unity) val = val / tot0;
            } else {
                const double src_x = ((double)(x + 1) - 0.5) * scale + 0.5, src_y = ((double)(y + 1) - 0.5) * scale + 0.5;
                int x0 = (int)floor(src_x); if (x0 < 1) x0 = 1;
                int x1 = x0 + 1; if (x1 > a.r0) x1 = a.r0;                    // min(size(image)[1], x0 + 1)
                int y0 = (int)floor(src_y); if (y0 < 1) y0 = 1;
                int y1 = y0 + 1; if (y1 > a.c0) y1 = a.c0;                    // min(size(image)[2], y0 + 1)
                const double dx = src_x - x0, dy = src_y - y0;
                double i00 = ingest_load(a, star, y0 - 1, x0 - 1), i01 = ingest_load(a, star, y0 - 1, x1 - 1);
                double i10 = ingest_load(a, star, y1 - 1, x0 - 1), i11 = ingest_load(a, star, y1 - 1, x1 - 1);
                if (a.unity) { i00 = i00 / tot0; i01 = i01 / tot0; i10 = i10 / tot0; i11 = i11 / tot0; }
                val = (1 - dx) * (1 - dy) * i00 + dx * (1 - dy) * i01 + (1 - dx) * dy * i10 + dx * dy * i11;
            }
            val = nm(val);
            out[p] = val;
            part += ntz(val);
        }
        double v2[1] = {part};
        cta_sum<1, 8>(v2, sred, buf);
        for (long long p = threadIdx.x; p < npx; p += 256) out[p] = out[p] / v2[0];     // same thread wrote it
    }
}

int ingest_launch(const void *d_in, int f32, int r0, int c0, long long N, int unity, int new_dim, int lanczos_a,
                  double *d_out, double *d_tmp, cudaStream_t st) {
    IngestArgs a;
    a.in = d_in; a.N = N; a.f32 = f32; a.r0 = r0; a.c0 = c0; a.unity = unity;
    a.resize = (new_dim != r0); a.nd = new_dim;
    a.out = (a.resize && lanczos_a > 0) ? d_tmp : d_out;
    long long grid = (long long)sm_count() * 8;
    if (grid > N) grid = N;
    ingest_kernel<<<(unsigned)grid, 256, 0, st>>>(a);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    if (a.resize && lanczos_a > 0) return smooth_launch(d_tmp, new_dim, new_dim, N, lanczos_a, 0, d_out, st);
    return SHOPT_OK;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int shopt_ingest_batch_dev(const void *d_vignets, int dtype_f32, int r0, int c0, int64_t N, int unity_sum, int new_dim,
                           int lanczos_a, double *d_stamps_out, void *stream) {
    if (!d_vignets || !d_stamps_out || r0 < 1 || c0 < 1 || N < 0 || new_dim < 1 || lanczos_a > 7) return SHOPT_E_BADARG;
    if (new_dim != r0 && r0 != c0) return SHOPT_E_BADARG;             // the reference resizes square stamps only
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (new_dim == r0 || lanczos_a <= 0)
        return ingest_launch(d_vignets, dtype_f32, r0, c0, N, unity_sum, new_dim, lanczos_a, d_stamps_out, nullptr, st);
    // resized stamps are smoothed from a scratch copy, a chunk of stars at a time
    const size_t px = (size_t)new_dim * new_dim, in_px = (size_t)r0 * c0, esz = dtype_f32 ? 4 : 8;
    const int64_t chunk = std::min<int64_t>(N, std::max<int64_t>(1, (int64_t)((size_t)1 << 30) / (int64_t)(px * 8)));
    TRY(ctx->need_scratch((size_t)chunk * px * 8));
    for (int64_t first = 0; first < N; first += chunk) {
        const int64_t cnt = std::min<int64_t>(chunk, N - first);
        TRY(ingest_launch((const char *)d_vignets + (size_t)first * in_px * esz, dtype_f32, r0, c0, cnt, unity_sum, new_dim,
                          lanczos_a, d_stamps_out + (size_t)first * px, (double *)ctx->scratch, st));
    }
    return SHOPT_OK;
}

int shopt_ingest_batch(const void *vignets, int dtype_f32, int r0, int c0, int64_t N, int unity_sum, int new_dim,
                       int lanczos_a, double *stamps_out) {
    if (!vignets || !stamps_out || r0 < 1 || c0 < 1 || N < 0 || new_dim < 1) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (N == 0) return SHOPT_OK;
    const size_t in_bytes = (size_t)N * r0 * c0 * (dtype_f32 ? 4 : 8);
    const size_t out_bytes = (size_t)N * (new_dim != r0 ? (size_t)new_dim * new_dim : (size_t)r0 * c0) * 8;
    DevBuf din, dout;
    TRY(din.alloc(in_bytes)); TRY(dout.alloc(out_bytes));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(din.p, vignets, in_bytes, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    TRY(shopt_ingest_batch_dev(din.p, dtype_f32, r0, c0, N, unity_sum, new_dim, lanczos_a, dout.as<double>(), st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(stamps_out, dout.p, out_bytes, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

// outliers_filter (outliers.jl:16-34): q1 = quantile(snr, k) on the Float32 SNR_WIN column (Statistics' default,
// type 7: h = (n - 1) k + 1, a + gamma (b - a) with (b - a) formed in Float32), indices with snr < q1.
int64_t shopt_snr_cut(const float *snr, int64_t N, double k, int64_t *idx_out, double *q1_out) {
    if (!snr || !idx_out || N < 1 || !(k >= 0.0 && k <= 1.0)) return SHOPT_E_BADARG;
    std::vector<float> v(snr, snr + N);
    for (float x : v) if (std::isnan(x)) return SHOPT_E_BADARG;       // Statistics.quantile raises on NaN
    std::sort(v.begin(), v.end());
    const double aleph = (double)N * k + (1.0 - k);                    // n p + m, m = alpha + p (1 - alpha - beta), alpha = beta = 1
    long long j = (long long)std::trunc(aleph);
    j = std::min<long long>(std::max<long long>(j, 1), std::max<long long>(N - 1, 1));
    double gamma = aleph - (double)j;
    gamma = std::min(std::max(gamma, 0.0), 1.0);
    const float a = v[(size_t)(j - 1)], b = (N > 1) ? v[(size_t)j] : v[0];
    const double q1 = (std::isfinite(a) && std::isfinite(b)) ? (double)a + gamma * (double)(float)(b - a)
                                                             : (1.0 - gamma) * (double)a + gamma * (double)b;
    if (q1_out) *q1_out = q1;
    int64_t m = 0;
    for (int64_t i = 0; i < N; i++) if ((double)snr[i] < q1) idx_out[m++] = i;
    return m;
}

}  // extern "C"

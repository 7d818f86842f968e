// api_polyfit.cu -- C ABI of the interpolation step (include/shopt_b200.h section 3 and the consumer p(u,v)).
#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

#include "api_ctx.h"
#include "polyfit_factor.h"

using namespace shopt;

extern "C" {

int64_t shopt_polyfit_acc_len(int r, int c, int degree) {
    if (r < 1 || c < 1 || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    const int64_t K = (degree + 1) * (degree + 2) / 2;
    return K * K + K * (int64_t)r * c;
}

int shopt_polyfit_centre_dev(const double *d_u, const double *d_v, int64_t n, double *d_centre, void *stream) {
    if (!d_u || !d_v || !d_centre || n < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    TRY(ctx->need_arena2(256));
    return uv_centre_launch(d_u, d_v, n, (double *)ctx->arena2, d_centre, Exchange(), (cudaStream_t)stream);
}

int shopt_polyfit_normal_dev(const double *d_pgf, int r, int c, const double *d_u, const double *d_v,
                             const int64_t *d_idx, int64_t n, int degree, const double *d_centre, double *d_acc, void *stream) {
    if (!d_pgf || !d_u || !d_v || !d_acc || r < 1 || c < 1 || n < 0 || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    const long long rc = (long long)r * c;
    TRY(ctx->need_scratch(polyfit_scratch_bytes(rc, n, degree)));
    StarList sl; sl.idx = (const long long *)d_idx; sl.n = n;
    return polyfit_normal_launch(d_pgf, rc, d_u, d_v, sl, degree, d_centre, d_acc, (double *)ctx->scratch, ctx->scratch_bytes,
                                 (cudaStream_t)stream);
}

namespace shopt {
__global__ void factor_head_kernel(const PolyFactor *F, int *info) { info[0] = F->rank; info[1] = F->info; }
}

int shopt_polyfit_solve_dev(const double *d_acc, int r, int c, int degree, const double *d_centre, int64_t n_rows,
                            double *d_poly, int32_t *d_info, void *stream) {
    if (!d_acc || !d_poly || r < 1 || c < 1 || degree < 0 || degree > 8 || n_rows < 0) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    TRY(ctx->need_arena2(sizeof(PolyFactor) + 512));
    PolyFactor *F = reinterpret_cast<PolyFactor *>((char *)ctx->arena2 + 256);
    cudaStream_t st = (cudaStream_t)stream;
    TRY(polyfit_factor_launch(d_acc, degree, d_centre, nullptr, n_rows, 0, F, st));
    TRY(polyfit_solve_launch(d_acc, (long long)r * c, degree, F, nullptr, nullptr, d_poly, st));
    if (d_info) { factor_head_kernel<<<1, 1, 0, st>>>(F, d_info); count_launches(1); }
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

int shopt_polyfit_factor_host(const double *G, int degree, const double *centre, int64_t n_rows, const double *b, int nrhs,
                              double *coef, int32_t *info_out) {
    if (!G || degree < 0 || degree > 8 || n_rows < 0 || nrhs < 0 || (nrhs > 0 && (!b || !coef))) return SHOPT_E_BADARG;
    const int K = (degree + 1) * (degree + 2) / 2;
    std::vector<double> work((size_t)7 * K * K + 2 * K);
    std::vector<PolyFactor> F(1);
    polyfit_factor(degree, G, centre, n_rows, 0, work.data(), F.data());
    for (int j = 0; j < nrhs; j++) {
        double bj[KMAX], cj[KMAX];
        for (int k = 0; k < K; k++) bj[k] = b[(size_t)k * nrhs + j];
        polyfit_apply(F.data(), bj, cj);
        for (int k = 0; k < K; k++) coef[(size_t)k * nrhs + j] = cj[k];
    }
    if (info_out) *info_out = F[0].info;
    return F[0].rank;
}

int shopt_polyfit_mse_dev(const double *d_pgf, int r, int c, const double *d_u, const double *d_v, const int64_t *d_idx,
                          int64_t n, int degree, const double *d_poly, double *d_mse, void *stream) {
    if (!d_pgf || !d_u || !d_v || !d_poly || !d_mse || r < 1 || c < 1 || n < 0 || degree < 0 || degree > 8)
        return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    const long long rc = (long long)r * c;
    TRY(ctx->need_scratch(polyfit_scratch_bytes(rc, n, degree)));
    StarList sl; sl.idx = (const long long *)d_idx; sl.n = n;
    return polyfit_mse_launch(d_pgf, rc, d_u, d_v, sl, degree, d_poly, d_mse, (double *)ctx->scratch, ctx->scratch_bytes,
                              (cudaStream_t)stream);
}

// shopt.jl:513-520: sort (index, error) by error, rev = true (stable: ties keep the lower index first; a NaN error is
// the largest under isless), take the first ceil(0.10 n).  sigma_clip: shopt.jl:522-532, err > mean + 3 std.
int64_t shopt_worst_10_percent(const double *err, int64_t n, int sigma_clip, int64_t *idx_out) {
    if (n < 0 || (n > 0 && (!err || !idx_out))) return SHOPT_E_BADARG;
    if (n == 0) return 0;
    if (sigma_clip) {
        // Julia's mean() is a pairwise sum over blocks of 1024
        struct PS { static double run(const double *a, int64_t m) {
            if (m <= 1024) { double s = 0.0; for (int64_t i = 0; i < m; i++) s += a[i]; return s; }
            const int64_t h = m >> 1; return run(a, h) + run(a + h, m - h); } };
        const double mean = PS::run(err, n) / (double)n;
        double ss = 0.0;
        for (int64_t i = 0; i < n; i++) ss += (err[i] - mean) * (err[i] - mean);
        const double thr = mean + 3 * std::sqrt(ss / (double)(n - 1));
        int64_t m = 0;
        for (int64_t i = 0; i < n; i++) if (err[i] > thr) idx_out[m++] = i;
        return m;
    }
    // The stable descending sort keeps ties in index order, i.e. it sorts by the TOTAL order (error descending with NaN
    // first, index ascending); its first m entries are therefore the m smallest under that order: select, then sort m.
    std::vector<int64_t> order((size_t)n);
    std::iota(order.begin(), order.end(), (int64_t)0);
    auto before = [err](int64_t a, int64_t b) {
        const double x = err[a], y = err[b];          // "a before b" <=> isless(y, x), ties by index
        const bool nx = std::isnan(x), ny = std::isnan(y);
        if (nx || ny) return (nx && ny) ? a < b : nx;
        if (x != y) return y < x;
        return a < b;
    };
    {
        const int64_t m = (int64_t)std::ceil(0.10 * (double)n);
        if (m < n) std::nth_element(order.begin(), order.begin() + m, order.end(), before);
        std::sort(order.begin(), order.begin() + m, before);
    }
    const int64_t m = (int64_t)std::ceil(0.10 * (double)n);
    for (int64_t i = 0; i < m; i++) idx_out[i] = order[(size_t)i];
    return m;
}

// shopt.jl:492-577 on one GPU, host buffers
int shopt_polyfit(const double *pgf, int r, int c, int64_t N, const double *u, const double *v, const int64_t *train_idx,
                  int64_t n_train, int degree, int iterations, int median_constraint, int sigma_clip, double *poly_out,
                  double *mse_out, int64_t *kept_idx, int64_t *n_kept, int64_t *n_per_round) {
    if (!pgf || !u || !v || !poly_out || r < 1 || c < 1 || N < 0 || n_train < 0 || (n_train > 0 && !train_idx))
        return SHOPT_E_BADARG;
    if (degree < 0 || degree > 8 || iterations < 0) return SHOPT_E_BADARG;
    for (int64_t t = 0; t < n_train; t++) if (train_idx[t] < 0 || train_idx[t] >= N) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    const long long rc = (long long)r * c;
    const int K = (degree + 1) * (degree + 2) / 2;
    const size_t n0 = (size_t)n_train;
    // only the training stars travel to the device, compacted in training order
    DevBuf dpgf, du, dv, didx, dacc, dpoly, dmse, dinfo, dmed, dfac, drows;
    if (median_constraint) TRY(dmed.alloc((size_t)rc * 8));
    TRY(dfac.alloc(2 * sizeof(PolyFactor) + 1024)); TRY(drows.alloc((size_t)4 * rc * 8 + 64));
    PolyFactor *F = dfac.as<PolyFactor>(), *Fred = F + 1;
    double *dctr = reinterpret_cast<double *>(Fred + 1), *dsums = dctr + 8;
    TRY(dpgf.alloc(std::max<size_t>(n0, 1) * rc * 8)); TRY(du.alloc(n0 * 8)); TRY(dv.alloc(n0 * 8)); TRY(didx.alloc(n0 * 8));
    TRY(dacc.alloc((size_t)shopt_polyfit_acc_len(r, c, degree) * 8)); TRY(dpoly.alloc((size_t)K * rc * 8));
    TRY(dmse.alloc(n0 * 8)); TRY(dinfo.alloc(16));
    TRY(ctx->need_scratch(polyfit_scratch_bytes(rc, n_train, degree)));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    std::vector<double> tu(n0), tv(n0);
    for (size_t t = 0; t < n0; t++) {
        tu[t] = u[train_idx[t]]; tv[t] = v[train_idx[t]];
        SHOPT_CUDA_TRY(cudaMemcpyAsync(dpgf.as<double>() + t * rc, pgf + (size_t)train_idx[t] * rc, (size_t)rc * 8,
                                       cudaMemcpyHostToDevice, st));
    }
    SHOPT_CUDA_TRY(cudaMemcpyAsync(du.p, tu.data(), n0 * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(dv.p, tv.data(), n0 * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    TRY(uv_centre_launch(du.as<double>(), dv.as<double>(), (long long)n0, dsums, dctr, Exchange(), st));
    std::vector<int64_t> cur(n0);             // positions in the compacted device batch
    std::iota(cur.begin(), cur.end(), (int64_t)0);
    std::vector<double> err(n0);
    std::vector<int64_t> bad(std::max<size_t>(n0, 1));
    // ones(r, c, K) (shopt.jl:492) is what survives zero rounds
    if (iterations == 0) for (size_t i = 0; i < (size_t)K * rc; i++) poly_out[i] = 1.0;
    for (int loop = 1; loop <= iterations; loop++) {
        const int64_t n = (int64_t)cur.size();
        if (n_per_round) n_per_round[loop - 1] = n;
        SHOPT_CUDA_TRY(cudaMemcpyAsync(didx.p, cur.data(), (size_t)n * 8, cudaMemcpyHostToDevice, st));
        StarList sl; sl.idx = didx.as<long long>(); sl.n = n;
        if (n < 5) {                           // interpolate.jl:142-145
            for (int64_t t = 0; t < n; t++)
                SHOPT_CUDA_TRY(cudaMemcpyAsync(drows.as<double>() + (size_t)t * rc, dpgf.as<double>() + (size_t)cur[(size_t)t] * rc,
                                               (size_t)rc * 8, cudaMemcpyDeviceToDevice, st));
            TRY(small_median_launch(drows.as<double>(), rc, (int)n, degree, dpoly.as<double>(), st));
        } else {
            TRY(polyfit_normal_launch(dpgf.as<double>(), rc, du.as<double>(), dv.as<double>(), sl, degree, dctr,
                                      dacc.as<double>(), (double *)ctx->scratch, ctx->scratch_bytes, st));
            if (!median_constraint)            // NaN values -> median before the solve (interpolate.jl:126-130)
                TRY(polyfit_median_launch(dpgf.as<double>(), rc, du.as<double>(), dv.as<double>(), sl, degree, dctr, 2,
                                          dacc.as<double>(), nullptr, st));
            TRY(polyfit_factor_launch(dacc.as<double>(), degree, dctr, nullptr, n, 0, F, st));
            if (median_constraint) {           // interpolate.jl:118-124, 139-141
                TRY(polyfit_factor_launch(dacc.as<double>(), degree, dctr, nullptr, n, 1, Fred, st));
                TRY(polyfit_median_launch(dpgf.as<double>(), rc, du.as<double>(), dv.as<double>(), sl, degree, dctr, 1,
                                          nullptr, dmed.as<double>(), st));
                TRY(polyfit_solve_launch(dacc.as<double>(), rc, degree, F, Fred, dmed.as<double>(), dpoly.as<double>(), st));
            } else {
                TRY(polyfit_solve_launch(dacc.as<double>(), rc, degree, F, nullptr, nullptr, dpoly.as<double>(), st));
            }
        }
        TRY(polyfit_mse_launch(dpgf.as<double>(), rc, du.as<double>(), dv.as<double>(), sl, degree, dpoly.as<double>(),
                               dmse.as<double>(), (double *)ctx->scratch, ctx->scratch_bytes, st));
        SHOPT_CUDA_TRY(cudaMemcpyAsync(err.data(), dmse.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
        SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
        if (loop == iterations) {
            if (mse_out) std::copy(err.begin(), err.begin() + n, mse_out);
        } else {                               // shopt.jl:557-575
            const int64_t nbad = shopt_worst_10_percent(err.data(), n, sigma_clip, bad.data());
            std::vector<char> isbad((size_t)n, 0);
            for (int64_t b = 0; b < nbad; b++) isbad[(size_t)bad[b]] = 1;
            size_t m = 0;
            for (int64_t t = 0; t < n; t++) if (!isbad[(size_t)t]) cur[m++] = cur[(size_t)t];
            cur.resize(m);
        }
    }
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    if (iterations > 0) SHOPT_CUDA_TRY(cudaMemcpyAsync(poly_out, dpoly.p, (size_t)K * rc * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    TRY(finish_timing(ctx));
    if (kept_idx) for (size_t t = 0; t < cur.size(); t++) kept_idx[t] = train_idx[cur[t]];
    if (n_kept) *n_kept = (int64_t)cur.size();
    return SHOPT_OK;
}

int shopt_psf_at_batch_dev(const double *d_poly, int r, int c, int degree, const double *d_u, const double *d_v, int64_t M,
                           double *d_stars_out, void *stream) {
    if (!d_poly || !d_u || !d_v || !d_stars_out || r < 1 || c < 1 || M < 0 || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    return psf_at_launch(d_poly, (long long)r * c, degree, d_u, d_v, M, d_stars_out, (cudaStream_t)stream);
}

int shopt_psf_at_batch(const double *poly, int r, int c, int degree, const double *u, const double *v, int64_t M,
                       double *stars_out) {
    if (!poly || !u || !v || !stars_out || r < 1 || c < 1 || M < 0 || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    if (M == 0) return SHOPT_OK;
    const size_t rc = (size_t)r * c, K = (size_t)(degree + 1) * (degree + 2) / 2, m = (size_t)M;
    DevBuf dpoly, duv, dout;
    TRY(dpoly.alloc(K * rc * 8)); TRY(duv.alloc(m * 16)); TRY(dout.alloc(m * rc * 8));
    cudaStream_t st = ctx->stream;
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(dpoly.p, poly, K * rc * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(duv.p, u, m * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(duv.as<double>() + m, v, m * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], st));
    TRY(psf_at_launch(dpoly.as<double>(), (long long)rc, degree, duv.as<double>(), duv.as<double>() + m, M,
                      dout.as<double>(), st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(stars_out, dout.p, m * rc * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    return finish_timing(ctx);
}

}  // extern "C"

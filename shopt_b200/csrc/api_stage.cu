// api_stage.cu -- the whole per-star fitting stage as one call (include/shopt_b200.h, shopt_stage_run*): the host
// control flow of shopt/shopt.jl:130-184, 281-314, 357-417, 472-577 around the kernels, for one GPU or one shard of a
// multi-GPU run (the caller supplies the all-reduce).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

#include "api_ctx.h"

namespace shopt {

// flags[i] |= BLACKLISTED from the model s (shopt.jl:166-171); |= PCA_FAILED / LEARNED_OOB (shopt.jl:305-309, 384-386)
__global__ void flag_kernel(const double *__restrict__ s_model, const int *__restrict__ pca_status,
                            const double *__restrict__ s_learned, long long n, double lo, double hi, int *__restrict__ flags) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int f = 0;
    const double sm = s_model[i];
    if (sm < lo || sm > hi) f |= SHOPT_FLAG_BLACKLISTED;
    if (pca_status[i] == SHOPT_PCA_FAILED) f |= SHOPT_FLAG_PCA_FAILED;
    const double sl = s_learned[i];
    if (sl < lo || sl > hi) f |= SHOPT_FLAG_LEARNED_OOB;
    flags[i] = f;
}


// Normal equations of the cubic field fit (shopt.jl:422-466) over the surviving stars of a shard:
// part[block][0..100) = T'T, [100..130) = T'[s g1 g2], T = the reference's ten terms.  Fixed-order sums: deterministic.
constexpr int FIELD_LEN = 130, FIELD_CHUNK = 128;
__global__ void __launch_bounds__(160) field_normal_kernel(const double *__restrict__ u, const double *__restrict__ v,
                                                           const double *__restrict__ learned, const int *__restrict__ flags,
                                                           long long n, long long per_block, double *__restrict__ part) {
    __shared__ double st[FIELD_CHUNK][13];
    const long long t0 = (long long)blockIdx.x * per_block, t1 = (t0 + per_block < n) ? t0 + per_block : n;
    const int e = threadIdx.x;
    const int i = (e < 100) ? e / 10 : (e - 100) % 10, j = (e < 100) ? e % 10 : 10 + (e - 100) / 10;
    double acc = 0.0;
    for (long long base = t0; base < t1; base += FIELD_CHUNK) {
        const int m = (int)((t1 - base < FIELD_CHUNK) ? (t1 - base) : FIELD_CHUNK);
        __syncthreads();
        if (e < m) {
            const long long g = base + e;
            const bool live = (flags[g] & 7) == 0;
            const double uu = u[g], vv = v[g];
            const double t[13] = {uu * uu * uu, vv * vv * vv, uu * uu * vv, vv * vv * uu, uu * uu, vv * vv, uu * vv, uu, vv, 1.0,
                                  learned[g], learned[n + g], learned[2 * n + g]};
#pragma unroll
            for (int k = 0; k < 13; k++) st[e][k] = live ? t[k] : 0.0;
        }
        __syncthreads();
        if (e < FIELD_LEN)
            for (int s = 0; s < m; s++) acc = fma(st[s][i], st[s][j], acc);
    }
    if (e < FIELD_LEN) part[(size_t)blockIdx.x * FIELD_LEN + e] = acc;
}
__global__ void field_reduce_kernel(const double *__restrict__ part, int nblocks, double *__restrict__ out) {
    const int e = threadIdx.x;
    if (e >= FIELD_LEN) return;
    double s = 0.0;
    for (int b = 0; b < nblocks; b++) s += part[(size_t)b * FIELD_LEN + e];
    out[e] = s;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int64_t shopt_training_split(const int32_t *flags, int64_t n_global, const int64_t *perm, double training_ratio,
                             int64_t *train_out) {
    if (n_global < 0 || (n_global > 0 && (!flags || !perm || !train_out))) return SHOPT_E_BADARG;
    int64_t n_surv = 0;
    for (int64_t g = 0; g < n_global; g++) n_surv += ((flags[g] & 7) == 0);
    const int64_t n_train = (int64_t)std::nearbyint(training_ratio * (double)n_surv);   // round half to even
    int64_t m = 0;
    for (int64_t i = 0; i < n_global && m < n_train; i++) {
        const int64_t g = perm[i];
        if (g < 0 || g >= n_global) return SHOPT_E_BADARG;
        if ((flags[g] & 7) == 0) train_out[m++] = g;
    }
    return m;
}

}  // extern "C"

namespace shopt {

// Per-star phase on stars [first, first + count) of a device-resident shard: model fit, pixel-grid fit, learned fit
// (shopt.jl:130-164, 281-314, 357-397).  Enqueues three launches on `st`; d_tmp = 3 ints per star of the shard.
static int stage_per_star(const shopt_params *p, const shopt_stage_io *io, long long first, long long count, int *d_tmp,
                          Ctx *ctx, cudaStream_t st) {
    if (count <= 0) return SHOPT_OK;
    const int r = io->r, c = io->c;
    const long long n = io->n_local, rc = (long long)r * c;
    const size_t nn = (size_t)std::max<long long>(n, 1);
    int *d_fst = d_tmp + first, *d_fit = d_tmp + nn + first, *d_pst = d_tmp + 2 * nn + first;
    double *m = io->model, *l = io->learned;
    const double *stamps = io->stamps + first * rc;
    double *pgf = io->pgf + first * rc;
    TRY(fit_analytic_launch(stamps, r, c, count, p->n_a, io->x0_model + 3 * first, p->g_tol_model, p->max_iter, m + first,
                            m + n + first, m + 2 * n + first, d_fst, d_fit, ctx->next_counter(), st));
    if (p->mode == SHOPT_MODE_PCA) {
        TRY(pca_launch(stamps, r, c, count, p->pca_terms, p->pratio, p->lanczos, p->unity_sum, pgf,
                       io->pca_k ? io->pca_k + first : nullptr, d_pst, ctx->next_counter(), (double *)ctx->scratch,
                       ctx->scratch_bytes, st));
    } else {   // element-wise modes never fail a star (shopt.jl:194-224, 316-348)
        if (p->mode == SHOPT_MODE_SMOOTH && !p->unity_sum)
            TRY(smooth_launch(stamps, r, c, count, p->lanczos, 1, pgf, st));
        else if (p->mode == SHOPT_MODE_SMOOTH)
            TRY(elem_pgf_launch(stamps, nullptr, nullptr, 0.0, rc, count, 0, 1, pgf, st));
        else
            TRY(elem_pgf_launch(stamps, io->weights + first * rc, io->x0_pixels ? io->x0_pixels + first * rc : nullptr,
                                p->chisq_fill, rc, count, 1, p->unity_sum ? 1 : 0, pgf, st));
        SHOPT_CUDA_TRY(cudaMemsetAsync(d_pst, 0, (size_t)count * 4, st));
        if (io->pca_k) SHOPT_CUDA_TRY(cudaMemsetAsync(io->pca_k + first, 0, (size_t)count * 4, st));
    }
    TRY(fit_analytic_launch(pgf, r, c, count, p->n_a, io->x0_learned + 3 * first, p->g_tol_learned, p->max_iter, l + first,
                            l + n + first, l + 2 * n + first, d_fst, d_fit, ctx->next_counter(), st));
    return SHOPT_OK;
}

static int stage_check_args(const shopt_params *p, const shopt_stage_io *io) {
    if (!p || !io || !io->stamps || !io->u || !io->v || !io->x0_model || !io->x0_learned || !io->model || !io->learned ||
        !io->pgf || !io->flags || !io->poly)
        return SHOPT_E_BADARG;
    const int r = io->r, c = io->c;
    const long long n = io->n_local, NG = io->n_global, off = io->offset;
    if (r < 1 || c < 2 || n < 0 || NG < n || off < 0 || off + n > NG || (NG > 0 && !io->perm)) return SHOPT_E_BADARG;
    if (p->degree < 0 || p->degree > 8 || p->iterations < 0 || p->n_a < 1 || p->n_a > r || p->n_a > c) return SHOPT_E_BADARG;
    if (p->mode < SHOPT_MODE_PCA || p->mode > SHOPT_MODE_CHISQ) return SHOPT_E_BADARG;
    if (p->mode == SHOPT_MODE_CHISQ && !io->weights) return SHOPT_E_BADARG;
    if (p->mode == SHOPT_MODE_SMOOTH && p->lanczos > 7) return SHOPT_E_BADARG;
    return SHOPT_OK;
}
static size_t stage_scratch_bytes(const shopt_params *p, const shopt_stage_io *io) {
    return std::max(p->mode == SHOPT_MODE_PCA ? pca_scratch_bytes(io->r, io->c, p->pca_terms) : (size_t)0,
                    polyfit_scratch_bytes((long long)io->r * io->c, io->n_local, p->degree));
}

// SHOPT_TRACE=1: host wall-clock milestones of one stage call on stderr (tracing aid; off by default)
struct StageTrace {
    bool on = getenv("SHOPT_TRACE") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now(), last = t0;
    void mark(const char *what, cudaStream_t st) {
        if (!on) return;
        cudaStreamSynchronize(st);
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[shopt trace] %-28s +%8.3f ms  (total %8.3f ms)\n", what,
                std::chrono::duration<double, std::milli>(now - last).count(),
                std::chrono::duration<double, std::milli>(now - t0).count());
        last = now;
    }
};

// The stage on a device-resident shard.  d_tmp: 3 ints per star (fit status | fit iterations | pca status);
// per_star_done: the caller has already enqueued stage_per_star() for every star on `st` (the pipelined host path).
static int stage_run_impl(const shopt_params *p, shopt_stage_io *io, shopt_allreduce_fn comm, void *comm_user, void *stream,
                          int *d_tmp, bool per_star_done) {
    Ctx *ctx; TRY(get_ctx(&ctx));
    const int r = io->r, c = io->c;
    const long long n = io->n_local, NG = io->n_global, off = io->offset;
    cudaStream_t st = (cudaStream_t)stream;
    const long long rc = (long long)r * c;
    const int K = (p->degree + 1) * (p->degree + 2) / 2;
    const size_t nn = (size_t)std::max<long long>(n, 1);
    DevBuf dacc, didx, dmse, dkeep, dinfo, dmed;
    if (p->median_constraint && comm) return SHOPT_E_UNSUPPORTED;
    if (p->median_constraint) TRY(dmed.alloc((size_t)rc * 8));
    int *d_pst = d_tmp + 2 * nn;
    double *m = io->model, *l = io->learned;
    // ---- model fit, pixel-grid fit, learned fit (every star of the shard; flags decide who counts) ----
    if (!per_star_done) TRY(stage_per_star(p, io, 0, n, d_tmp, ctx, st));
    if (n > 0) flag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m, d_pst, l, n, p->s_lo, p->s_hi, io->flags);
    SHOPT_CUDA_TRY(cudaGetLastError());
    StageTrace tr;
    tr.mark("per-star kernels + flags", st);
    std::vector<int> flags((size_t)n);
    SHOPT_CUDA_TRY(cudaMemcpyAsync(flags.data(), io->flags, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));

    // ---- survivors of the whole catalogue (a zero-padded gather through the all-reduce) ----
    std::vector<int> gflag((size_t)NG, 0);
    for (long long i = 0; i < n; i++) gflag[(size_t)(off + i)] = flags[(size_t)i] + 32;   // +32 marks "reported"
    if (comm) {
        TRY(dkeep.alloc((size_t)std::max<long long>(NG, 1) * 4));
        SHOPT_CUDA_TRY(cudaMemcpyAsync(dkeep.p, gflag.data(), (size_t)NG * 4, cudaMemcpyHostToDevice, st));
        if (comm(dkeep.p, NG, 1, stream, comm_user) != 0) { set_last_error_text("all-reduce callback failed"); return SHOPT_E_CUDA; }
        SHOPT_CUDA_TRY(cudaMemcpyAsync(gflag.data(), dkeep.p, (size_t)NG * 4, cudaMemcpyDeviceToHost, st));
        SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    }
    tr.mark("flag gather", st);
    long long n_black = 0, n_failed = 0, n_surv = 0;
    for (long long g = 0; g < NG; g++) {
        const int f = gflag[(size_t)g] & 31;
        if (f & SHOPT_FLAG_BLACKLISTED) n_black++;
        else if (f & (SHOPT_FLAG_PCA_FAILED | SHOPT_FLAG_LEARNED_OOB)) n_failed++;
        else n_surv++;
    }
    // ---- s(u,v), g1(u,v), g2(u,v) over the survivors' learned fits (shopt.jl:422-466): normal equations of the cubic
    //      accumulated on the device (additive over shards -> one 130-double all-reduce), solved on the host ----
    {
        DevBuf dfield;
        const int nb = (int)std::min<long long>(64, std::max<long long>(1, (n + FIELD_CHUNK - 1) / FIELD_CHUNK));
        TRY(dfield.alloc((size_t)(nb + 1) * FIELD_LEN * 8));
        double *d_part = dfield.as<double>() + FIELD_LEN, *d_sum = dfield.as<double>();
        const long long per_block = (n + nb - 1) / nb;
        field_normal_kernel<<<nb, 160, 0, st>>>(io->u, io->v, l, io->flags, n, per_block, d_part);
        field_reduce_kernel<<<1, 160, 0, st>>>(d_part, nb, d_sum);
        SHOPT_CUDA_TRY(cudaGetLastError());
        if (comm && comm(d_sum, FIELD_LEN, 0, stream, comm_user) != 0) { set_last_error_text("all-reduce callback failed"); return SHOPT_E_CUDA; }
        double h[FIELD_LEN];
        SHOPT_CUDA_TRY(cudaMemcpyAsync(h, d_sum, sizeof h, cudaMemcpyDeviceToHost, st));
        SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
        io->field_rank = field_solve_normal(h, h + 100, 3, io->field);
    }
    tr.mark("field fit", st);
    // ---- training split: first round(training_ratio * n_surv) survivors in perm order (shopt.jl:472-478) ----
    for (auto &f : gflag) f &= 31;
    std::vector<int64_t> train((size_t)std::max<long long>(n_surv, 1));   // global star ids, training order
    const int64_t n_train = shopt_training_split(gflag.data(), NG, io->perm, p->training_ratio, train.data());
    if (n_train < 0) return (int)n_train;
    train.resize((size_t)n_train);
    for (int64_t g : train) if (g >= off && g < off + n) flags[(size_t)(g - off)] |= SHOPT_FLAG_TRAINING;

    tr.mark("training split", st);
    // ---- polyfit rounds (shopt.jl:534-577) ----
    const long long acc_len = (long long)K * K + (long long)K * rc;
    TRY(dacc.alloc((size_t)acc_len * 8)); TRY(didx.alloc((size_t)std::max<long long>(n, 1) * 8));
    TRY(dmse.alloc((size_t)std::max<size_t>(train.size(), 1) * 8)); TRY(dinfo.alloc(16));
    std::vector<long long> cur(train.begin(), train.end());   // alive training stars (global ids) in training order
    std::vector<long long> lidx, lpos;               // this shard's members: local index, position in `cur`
    std::vector<double> err, lerr;
    std::vector<int64_t> bad;
    if (p->iterations == 0) {                        // ones(r, c, K), shopt.jl:492
        std::vector<double> ones((size_t)K * rc, 1.0);
        SHOPT_CUDA_TRY(cudaMemcpyAsync(io->poly, ones.data(), ones.size() * 8, cudaMemcpyHostToDevice, st));
        SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    }
    for (int loop = 1; loop <= p->iterations; loop++) {
        const long long nc = (long long)cur.size();
        if (io->n_per_round) io->n_per_round[loop - 1] = nc;
        lidx.clear(); lpos.clear();
        for (long long t = 0; t < nc; t++)
            if (cur[(size_t)t] >= off && cur[(size_t)t] < off + n) { lidx.push_back(cur[(size_t)t] - off); lpos.push_back(t); }
        const long long nl = (long long)lidx.size();
        SHOPT_CUDA_TRY(cudaMemcpyAsync(didx.p, lidx.data(), (size_t)nl * 8, cudaMemcpyHostToDevice, st));
        if (nc < 5) {                                // interpolate.jl:142-145
            if (comm) return SHOPT_E_UNSUPPORTED;    // fewer than 5 training stars across a sharded run
            TRY(small_median_launch(io->pgf, rc, didx.as<long long>(), (int)nl, p->degree, io->poly, st));
        } else {
            TRY(polyfit_normal_launch(io->pgf, rc, io->u, io->v, didx.as<long long>(), nl, p->degree, dacc.as<double>(),
                                      (double *)ctx->scratch, ctx->scratch_bytes, st));
            if (comm && comm(dacc.p, acc_len, 0, stream, comm_user) != 0) {
                set_last_error_text("all-reduce callback failed"); return SHOPT_E_CUDA;
            }
            // per-pixel medians over the training stars (median_constraint; NaN pixel-grid values) are selected on one
            // GPU; a sharded run would need a histogram exchange per radix pass and is refused above
            if (!comm && !p->median_constraint)
                TRY(polyfit_median_launch(io->pgf, rc, io->u, io->v, didx.as<long long>(), nl, p->degree, 2,
                                          dacc.as<double>(), nullptr, st));
            tr.mark("  normal + all-reduce", st);
            TRY(polyfit_solve_launch(dacc.as<double>(), rc, p->degree, io->poly, dinfo.as<int>(), st));
            if (p->median_constraint) {
                TRY(polyfit_median_launch(io->pgf, rc, io->u, io->v, didx.as<long long>(), nl, p->degree, 1, nullptr,
                                          dmed.as<double>(), st));
                TRY(polyfit_set_coef0_launch(dmed.as<double>(), rc, io->poly, st));
            }
        }
        TRY(polyfit_mse_launch(io->pgf, rc, io->u, io->v, didx.as<long long>(), nl, p->degree, io->poly, dmse.as<double>(),
                               (double *)ctx->scratch, ctx->scratch_bytes, st));
        lerr.resize((size_t)nl);
        SHOPT_CUDA_TRY(cudaMemcpyAsync(lerr.data(), dmse.p, (size_t)nl * 8, cudaMemcpyDeviceToHost, st));
        SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
        tr.mark("  solve + mse", st);
        err.assign((size_t)nc, 0.0);
        for (long long i = 0; i < nl; i++) err[(size_t)lpos[(size_t)i]] = lerr[(size_t)i];
        if (comm && nc > 0) {                        // zero-padded gather of the per-star errors
            SHOPT_CUDA_TRY(cudaMemcpyAsync(dmse.p, err.data(), (size_t)nc * 8, cudaMemcpyHostToDevice, st));
            if (comm(dmse.p, nc, 0, stream, comm_user) != 0) { set_last_error_text("all-reduce callback failed"); return SHOPT_E_CUDA; }
            SHOPT_CUDA_TRY(cudaMemcpyAsync(err.data(), dmse.p, (size_t)nc * 8, cudaMemcpyDeviceToHost, st));
            SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
        }
        tr.mark("  error gather", st);
        if (loop < p->iterations) {                  // shopt.jl:557-575; the last round's selection is not applied
            bad.resize((size_t)std::max<long long>(nc, 1));
            const int64_t nbad = shopt_worst_10_percent(err.data(), nc, p->sigma_clip, bad.data());
            std::vector<char> isbad((size_t)nc, 0);
            for (int64_t b = 0; b < nbad; b++) isbad[(size_t)bad[(size_t)b]] = 1;
            size_t mkeep = 0;
            for (long long t = 0; t < nc; t++) {
                const long long g = cur[(size_t)t];
                if (isbad[(size_t)t]) { if (g >= off && g < off + n) flags[(size_t)(g - off)] |= SHOPT_FLAG_REJECTED; }
                else cur[mkeep++] = g;
            }
            cur.resize(mkeep);
        }
    }
    tr.mark("rounds done", st);
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->flags, flags.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    io->counts[0] = n_black; io->counts[1] = n_failed; io->counts[2] = (int64_t)train.size(); io->counts[3] = (int64_t)cur.size();
    return SHOPT_OK;
}

}  // namespace shopt

extern "C" {

int shopt_stage_run_dev(const shopt_params *p, shopt_stage_io *io, shopt_allreduce_fn comm, void *comm_user, void *stream) {
    TRY(stage_check_args(p, io));
    Ctx *ctx; TRY(get_ctx(&ctx));
    DevBuf dint;
    TRY(dint.alloc((size_t)std::max<long long>(io->n_local, 1) * 12));
    TRY(ctx->need_scratch(stage_scratch_bytes(p, io)));
    return stage_run_impl(p, io, comm, comm_user, stream, dint.as<int>(), false);
}

// Host buffers, one GPU.  The stamps go up in chunks on a copy stream while the per-star kernels of the chunks that
// have landed run on the compute stream, so the H2D transfer (13.7 GB at 100k x 131 x 131) hides behind the fits.
int shopt_stage_run(const shopt_params *p, shopt_stage_io *io) { return shopt_stage_run_comm(p, io, nullptr, nullptr); }

int shopt_stage_run_comm(const shopt_params *p, shopt_stage_io *io, shopt_allreduce_fn comm, void *comm_user) {
    if (!p || !io || !io->stamps || !io->u || !io->v || !io->x0_model || !io->x0_learned || !io->model || !io->learned ||
        !io->flags || !io->poly)
        return SHOPT_E_BADARG;
    if (io->r < 1 || io->c < 2 || io->n_local < 0) return SHOPT_E_BADARG;
    if (!comm && (io->n_global != io->n_local || io->offset != 0)) return SHOPT_E_BADARG;
    Ctx *ctx; TRY(get_ctx(&ctx));
    const size_t n = (size_t)io->n_local, rc = (size_t)io->r * io->c, nn = std::max<size_t>(n, 1);
    const size_t K = (size_t)(p->degree + 1) * (p->degree + 2) / 2;
    DevBuf dsm, dout, dint, dpoly, dtmp;
    RawBuf dst, dpgf;
    TRY(ctx->need_cache(0, nn * rc * 8, &dst.p)); TRY(ctx->need_cache(1, nn * rc * 8, &dpgf.p));
    TRY(dsm.alloc(nn * 8 * 8)); TRY(dout.alloc(nn * 6 * 8));
    TRY(dint.alloc(nn * 8)); TRY(dpoly.alloc(K * rc * 8)); TRY(dtmp.alloc(nn * 12));
    DevBuf dw, dx0p;                                 // mode chisq only
    const bool chisq = p->mode == SHOPT_MODE_CHISQ;
    if (chisq && !io->weights) return SHOPT_E_BADARG;
    if (chisq) TRY(dw.alloc(nn * rc * 8));
    if (chisq && io->x0_pixels) TRY(dx0p.alloc(nn * rc * 8));
    cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
    double *small = dsm.as<double>();                // u | v | x0_model (3n) | x0_learned (3n)
    shopt_stage_io d = *io;
    d.stamps = dst.as<double>(); d.u = small; d.v = small + n; d.x0_model = small + 2 * n; d.x0_learned = small + 5 * n;
    d.model = dout.as<double>(); d.learned = dout.as<double>() + 3 * n; d.pgf = dpgf.as<double>();
    d.flags = dint.as<int32_t>(); d.pca_k = io->pca_k ? dint.as<int32_t>() + n : nullptr; d.poly = dpoly.as<double>();
    d.weights = chisq ? dw.as<double>() : nullptr; d.x0_pixels = (chisq && io->x0_pixels) ? dx0p.as<double>() : nullptr;
    TRY(stage_check_args(p, &d));
    TRY(ctx->need_scratch(stage_scratch_bytes(p, &d)));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaStreamWaitEvent(cs, ctx->ev[0], 0));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small, io->u, n * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small + n, io->v, n * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small + 2 * n, io->x0_model, n * 24, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small + 5 * n, io->x0_learned, n * 24, cudaMemcpyHostToDevice, st));
    // chunks: at least ~16 stars per resident CTA so the persistent kernels keep their tail short; at most NCHUNK_EV
    long long nchunks = (long long)n / (16LL * 3 * sm_count());
    if (nchunks > Ctx::NCHUNK_EV) nchunks = Ctx::NCHUNK_EV;
    if (nchunks < 1) nchunks = 1;
    const long long per = ((long long)n + nchunks - 1) / nchunks;
    for (long long ck = 0; ck < nchunks; ck++) {
        const long long first = ck * per, count = std::min<long long>(per, (long long)n - first);
        if (count <= 0) break;
        SHOPT_CUDA_TRY(cudaMemcpyAsync(dst.as<double>() + first * rc, io->stamps + first * rc, (size_t)count * rc * 8,
                                       cudaMemcpyHostToDevice, cs));
        if (chisq) SHOPT_CUDA_TRY(cudaMemcpyAsync(dw.as<double>() + first * rc, io->weights + first * rc,
                                                  (size_t)count * rc * 8, cudaMemcpyHostToDevice, cs));
        if (d.x0_pixels) SHOPT_CUDA_TRY(cudaMemcpyAsync(dx0p.as<double>() + first * rc, io->x0_pixels + first * rc,
                                                        (size_t)count * rc * 8, cudaMemcpyHostToDevice, cs));
        SHOPT_CUDA_TRY(cudaEventRecord(ctx->chunk_ev[ck], cs));
        SHOPT_CUDA_TRY(cudaStreamWaitEvent(st, ctx->chunk_ev[ck], 0));
        TRY(stage_per_star(p, &d, first, count, dtmp.as<int>(), ctx, st));
    }
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], cs));    // last stamp on the device
    TRY(stage_run_impl(p, &d, comm, comm_user, st, dtmp.as<int>(), true));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->model, d.model, n * 24, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->learned, d.learned, n * 24, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->flags, d.flags, n * 4, cudaMemcpyDeviceToHost, st));
    if (io->pca_k) SHOPT_CUDA_TRY(cudaMemcpyAsync(io->pca_k, d.pca_k, n * 4, cudaMemcpyDeviceToHost, st));
    if (io->pgf) SHOPT_CUDA_TRY(cudaMemcpyAsync(io->pgf, d.pgf, n * rc * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->poly, d.poly, K * rc * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    for (int i = 0; i < 4; i++) io->counts[i] = d.counts[i];
    for (int i = 0; i < 30; i++) io->field[i] = d.field[i];
    io->field_rank = d.field_rank;
    return finish_timing(ctx);
}

}  // extern "C"

// api_stage.cu -- the whole per-star fitting stage as one call (include/shopt_b200.h, shopt_stage_run*): the control
// flow of shopt/shopt.jl:130-184, 281-314, 357-417, 472-577 around the kernels, for one GPU or one shard of a sharded
// run.  The flow is ENQUEUE-ONLY: survivor counting, the training split, worst_10_percent and the compaction of the
// training list run on the device (select.cu); every buffer comes from the context's grow-only arena and results come
// back through pinned memory.  A call synchronises twice (once after the per-star phase, to learn the round sizes;
// once at the end) and allocates nothing after the first call of a given size.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "api_ctx.h"
#include "polyfit_factor.h"
#include "stage_state.h"

namespace shopt {

// Normal equations of the cubic field fit (shopt.jl:422-466) over the surviving stars of a shard, in the CENTRED basis
// and in polynomial_optimizer's term order (so that the polyfit factor / solve kernels serve it with rc = 3):
// part[block][0..100) = T'T, [100 + 3 k + j] = (T'[s g1 g2])[k][j].  Fixed-order sums: deterministic.
constexpr int FIELD_LEN = 130, FIELD_CHUNK = 128;
__global__ void __launch_bounds__(160) field_normal_kernel(const double *__restrict__ u, const double *__restrict__ v,
                                                           const double *__restrict__ learned, const int *__restrict__ flags,
                                                           const double *__restrict__ ctr, long long n, long long per_block,
                                                           double *__restrict__ part) {
    __shared__ double st[FIELD_CHUNK][13];
    const long long t0 = (long long)blockIdx.x * per_block, t1 = (t0 + per_block < n) ? t0 + per_block : n;
    const int e = threadIdx.x;
    const int i = (e < 100) ? e / 10 : (e - 100) / 3, j = (e < 100) ? e % 10 : 10 + (e - 100) % 3;
    double acc = 0.0;
    for (long long base = t0; base < t1; base += FIELD_CHUNK) {
        const int m = (int)((t1 - base < FIELD_CHUNK) ? (t1 - base) : FIELD_CHUNK);
        __syncthreads();
        if (e < m) {
            const long long g = base + e;
            const bool live = (flags[g] & 7) == 0;
            const double uu = (u[g] - ctr[0]) * ctr[1], vv = (v[g] - ctr[2]) * ctr[3];
            // p outer, q inner: 1, v, v^2, v^3, u, u v, u v^2, u^2, u^2 v, u^3
            const double t[13] = {1.0, vv, vv * vv, vv * vv * vv, uu, uu * vv, uu * vv * vv, uu * uu, uu * uu * vv, uu * uu * uu,
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
// polynomial_optimizer's order -> the reference's [u^3, v^3, u^2 v, v^2 u, u^2, v^2, u v, u, v, 1] (interpolate.jl:8);
// poly3[j + 3 k] = coefficient k of value j
__global__ void field_reorder_kernel(const double *__restrict__ poly3, double *__restrict__ field) {
    const int map[10] = {9, 3, 8, 6, 7, 2, 5, 4, 1, 0};
    const int e = threadIdx.x;
    if (e < 30) field[e] = poly3[(e / 10) + 3 * map[e % 10]];
}
// pixels whose V'P is NaN (a NaN pixel-grid value among the training stars): counted for the sharded path, which cannot
// repair them locally
__global__ void nan_count_kernel(const double *__restrict__ b0, long long rc, long long *__restrict__ count) {
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int c = __syncthreads_count(px < rc && isnan(b0[px]));
    if (threadIdx.x == 0 && c) atomicAdd((unsigned long long *)count, (unsigned long long)c);
}
__global__ void fill_kernel(double *__restrict__ p, long long n, double val) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = val;
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

// ---- the stage's device buffers, carved from the context's arena ------------------------------------------------------
struct StageBufs {
    int *tmp;                       // 5 n: model status | model iters | pca status | learned status | learned iters
    long long *perm;                // NG (null when the caller passed a device copy)
    unsigned *lbits, *gbits;        // W words each; gbits is followed by the 3 survivor counters (one all-reduce)
    int *counters, *bad_perm;
    StageState *state;
    long long *idx[2];              // training list (local star indices), ping-pong
    long long *rej_idx;             // the stars the last compaction dropped (n_rej of them): what the next round downdates by
    unsigned *pos[2];               // ... and each entry's position in the permutation (the tie-break of the selection)
    double *mse;
    unsigned *hist;                 // SEL_BINS
    int *blockcnt;
    long long *blockoff;
    double *acc, *rows, *med, *uvsums, *ctr, *clip;
    double *acc_raw, *acc_sub;      // the rounds' normal equations before any NaN repair, and the rejected stars' share of them
    void *shard_med;                // sharded runs: work space of the cross-shard per-pixel median (polyfit_median_sharded_launch)
    PolyFactor *fac, *fac_red, *fac_field;
    double *field_part, *field_acc, *field_poly, *field_out;
    int field_nb;
    size_t bytes;
};
static StageBufs stage_carve(void *base, const shopt_params *p, const shopt_stage_io *io, bool sharded) {
    StageBufs b{};
    Bump a(base);
    const size_t n = (size_t)std::max<long long>(io->n_local, 1), NG = (size_t)std::max<long long>(io->n_global, 1);
    const size_t rc = (size_t)io->r * io->c, W = (NG + 31) / 32;
    const size_t K = (size_t)(p->degree + 1) * (p->degree + 2) / 2;
    b.tmp = a.take<int>(5 * n);
    b.perm = io->perm_dev ? nullptr : a.take<long long>(NG);
    b.lbits = a.take<unsigned>(W + 8);
    b.gbits = sharded ? a.take<unsigned>(W + 8) : b.lbits;
    b.counters = reinterpret_cast<int *>(b.gbits ? b.gbits + W : nullptr);
    b.bad_perm = a.take<int>(4);
    b.state = a.take<StageState>(1);
    for (int i = 0; i < 2; i++) { b.idx[i] = a.take<long long>(n); b.pos[i] = a.take<unsigned>(n); }
    b.rej_idx = a.take<long long>(n);
    b.mse = a.take<double>(n);
    b.hist = a.take<unsigned>(SEL_BINS);
    const size_t nblk = (std::max(n, NG) + 1023) / 1024 + 1;
    b.blockcnt = a.take<int>(nblk);
    b.blockoff = a.take<long long>(nblk);
    b.acc = a.take<double>(K * K + K * rc);
    b.acc_raw = a.take<double>(K * K + K * rc); b.acc_sub = a.take<double>(K * K + K * rc);
    b.rows = a.take<double>(4 * rc);
    b.med = a.take<double>(rc);
    b.shard_med = sharded ? (void *)a.take<char>(polyfit_shard_median_bytes((long long)rc, p->degree)) : nullptr;
    b.uvsums = a.take<double>(8); b.ctr = a.take<double>(4); b.clip = a.take<double>(4);
    b.fac = a.take<PolyFactor>(1); b.fac_red = a.take<PolyFactor>(1); b.fac_field = a.take<PolyFactor>(1);
    b.field_nb = (int)std::min<long long>(64, std::max<long long>(1, (io->n_local + FIELD_CHUNK - 1) / FIELD_CHUNK));
    b.field_part = a.take<double>((size_t)b.field_nb * FIELD_LEN);
    b.field_acc = a.take<double>(FIELD_LEN); b.field_poly = a.take<double>(32); b.field_out = a.take<double>(32);
    b.bytes = a.off;
    return b;
}
// what comes back to the host, in pinned memory
struct StageResult {
    StageState state;
    int fac_head[4];        // K, rc_rank, rank, info of the last round's factorisation
    int field_head[4];
    double field[32];
    int bad_perm[4];
    long long n_nan;        // sharded runs: pixels whose all-reduced V'P is NaN this round
    long long n_after[16];  // sigma clipping: the whole catalogue's training count after each round's rejection
};

static int stage_check_args(const shopt_params *p, const shopt_stage_io *io) {
    if (!p || !io || !io->stamps || !io->u || !io->v || !io->x0_model || !io->x0_learned || !io->model || !io->learned ||
        !io->pgf || !io->flags || !io->poly)
        return SHOPT_E_BADARG;
    const int r = io->r, c = io->c;
    const long long n = io->n_local, NG = io->n_global, off = io->offset;
    if (r < 1 || c < 2 || n < 0 || NG < n || off < 0 || off + n > NG || (NG > 0 && !io->perm && !io->perm_dev)) return SHOPT_E_BADARG;
    if (NG >= (1LL << 32)) return SHOPT_E_UNSUPPORTED;                       // permutation positions are 32-bit keys
    if (p->degree < 0 || p->degree > 8 || p->iterations < 0 || p->n_a < 1 || p->n_a > r || p->n_a > c) return SHOPT_E_BADARG;
    if (p->iterations > 16 && p->sigma_clip) return SHOPT_E_UNSUPPORTED;
    if (p->mode < SHOPT_MODE_PCA || p->mode > SHOPT_MODE_CHISQ) return SHOPT_E_BADARG;
    if (p->profile < SHOPT_PROFILE_GAUSSIAN || p->profile > SHOPT_PROFILE_KOLMOGOROV) return SHOPT_E_BADARG;
    if (p->mode == SHOPT_MODE_CHISQ && !io->weights) return SHOPT_E_BADARG;
    if (p->mode == SHOPT_MODE_SMOOTH && p->lanczos > 7) return SHOPT_E_BADARG;
    return SHOPT_OK;
}
static size_t stage_scratch_bytes(const shopt_params *p, const shopt_stage_io *io) {
    return std::max(p->mode == SHOPT_MODE_PCA ? pca_scratch_bytes(io->r, io->c, p->pca_terms, io->n_local) : (size_t)0,
                    polyfit_scratch_bytes((long long)io->r * io->c, io->n_local, p->degree));
}

// First thing on the stream, before any per-star kernel: the permutation goes up (a copy from pageable memory blocks
// the host until the stream reaches it), the control state is zeroed, the (u, v) centre of the interpolation basis is
// computed (and all-reduced).
static int stage_begin(const shopt_params *p, const shopt_stage_io *io, const Exchange &ex, const StageBufs &b, cudaStream_t st) {
    (void)p;
    if (b.perm && io->n_global > 0)
        SHOPT_CUDA_TRY(cudaMemcpyAsync(b.perm, io->perm, (size_t)io->n_global * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemsetAsync(b.state, 0, sizeof(StageState), st));
    SHOPT_CUDA_TRY(cudaMemsetAsync(b.bad_perm, 0, 16, st));
    SHOPT_CUDA_TRY(cudaMemsetAsync(b.hist, 0, SEL_BINS * sizeof(unsigned), st));
    return uv_centre_launch(io->u, io->v, io->n_local, b.uvsums, b.ctr, ex, st);
}

// Per-star phase on stars [first, first + count) of a device-resident shard: model fit, pixel-grid fit, learned fit
// (shopt.jl:130-164, 281-314, 357-397).  Enqueues three launches on `st`.
static int stage_per_star(const shopt_params *p, const shopt_stage_io *io, long long first, long long count, int *d_tmp,
                          Ctx *ctx, cudaStream_t st) {
    if (count <= 0) return SHOPT_OK;
    const int r = io->r, c = io->c;
    const long long n = io->n_local, rc = (long long)r * c;
    const size_t nn = (size_t)std::max<long long>(n, 1);
    int *d_mst = d_tmp + first, *d_mit = d_tmp + nn + first, *d_pst = d_tmp + 2 * nn + first;
    int *d_lst = d_tmp + 3 * nn + first, *d_lit = d_tmp + 4 * nn + first;
    double *m = io->model, *l = io->learned;
    const double *stamps = io->stamps + first * rc;
    double *pgf = io->pgf + first * rc;
    TRY(fit_analytic_launch(stamps, r, c, count, p->n_a, io->x0_model + 3 * first, p->g_tol_model, p->max_iter, m + first,
                            m + n + first, m + 2 * n + first, d_mst, d_mit, ctx->next_counter(), st, nullptr, nullptr, p->profile));
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
                            l + n + first, l + 2 * n + first, d_lst, d_lit, ctx->next_counter(), st, nullptr, nullptr, p->profile));
    return SHOPT_OK;
}

// SHOPT_TRACE=1: host wall-clock milestones of one stage call on stderr (tracing aid; off by default; it adds syncs)
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

// The stage on a device-resident shard after stage_begin() (and, when per_star_done, after the caller has enqueued
// stage_per_star() for every star on `st`: the pipelined host path).
static int stage_run_impl(const shopt_params *p, shopt_stage_io *io, const Exchange &ex, cudaStream_t st, const StageBufs &b,
                          bool per_star_done) {
    Ctx *ctx; TRY(get_ctx(&ctx));
    const int r = io->r, c = io->c;
    const long long n = io->n_local, NG = io->n_global, off = io->offset;
    const long long rc = (long long)r * c;
    const int K = (p->degree + 1) * (p->degree + 2) / 2;
    const size_t nn = (size_t)std::max<long long>(n, 1);
    const long long W = (NG + 31) >> 5;
    const long long *perm = io->perm_dev ? (const long long *)io->perm_dev : b.perm;
    StageResult *res = static_cast<StageResult *>(ctx->pinned);
    memset(res, 0, sizeof *res);
    double *m = io->model, *l = io->learned;
    StageTrace tr;
    // ---- model fit, pixel-grid fit, learned fit (every star of the shard; flags decide who counts) ----
    if (!per_star_done) TRY(stage_per_star(p, io, 0, n, b.tmp, ctx, st));
    TRY(flag_launch(m, b.tmp + 2 * nn, l, b.tmp + 3 * nn, n, p->s_lo, p->s_hi, io->flags, b.counters, st));
    tr.mark("per-star kernels + flags", st);
    // ---- survivors along the permutation, training split (shopt.jl:472-478), this shard's training list ----
    TRY(surv_bits_launch(perm, NG, off, n, io->flags, b.lbits, b.gbits, b.bad_perm, st));
    if (ex.on()) TRY(ex.sum(b.gbits, W + 3, 1));
    TRY(split_launch(b.gbits, NG, p->training_ratio, b.counters, b.state, st));
    TRY(training_list_launch(b.lbits, perm, NG, off, b.state, b.idx[0], b.pos[0], io->flags, b.blockcnt, b.blockoff, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(&res->state, b.state, sizeof(StageState), cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(res->bad_perm, b.bad_perm, 16, cudaMemcpyDeviceToHost, st));
    // ---- s(u,v), g1(u,v), g2(u,v) over the survivors' learned fits (shopt.jl:422-466): the cubic's normal equations
    //      (additive over shards: one 130-double all-reduce) through the same factor / solve kernels as the pixels ----
    {
        const long long per_block = (n + b.field_nb - 1) / b.field_nb;
        field_normal_kernel<<<b.field_nb, 160, 0, st>>>(io->u, io->v, l, io->flags, b.ctr, n, per_block, b.field_part);
        field_reduce_kernel<<<1, 160, 0, st>>>(b.field_part, b.field_nb, b.field_acc);
        count_launches(2);
        SHOPT_CUDA_TRY(cudaGetLastError());
        TRY(ex.sum(b.field_acc, FIELD_LEN, 0));
        TRY(polyfit_factor_launch(b.field_acc, 3, b.ctr, &b.state->n_surv, 0, 0, b.fac_field, st));
        TRY(polyfit_solve_launch(b.field_acc, 3, 3, b.fac_field, nullptr, nullptr, b.field_poly, st));
        field_reorder_kernel<<<1, 32, 0, st>>>(b.field_poly, b.field_out);
        count_launches(1);
        SHOPT_CUDA_TRY(cudaMemcpyAsync(res->field, b.field_out, 30 * 8, cudaMemcpyDeviceToHost, st));
        SHOPT_CUDA_TRY(cudaMemcpyAsync(res->field_head, b.fac_field, 16, cudaMemcpyDeviceToHost, st));
    }
    // ---- sync 1 of 2: the host learns the survivor counts, i.e. the size of every round ----
    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    tr.mark("split + field fit", st);
    if (res->bad_perm[0]) { set_last_error_text("perm holds an index outside [0, n_global)"); return SHOPT_E_BADARG; }
    const long long n_train = res->state.n_train;
    io->counts[0] = res->state.n_black; io->counts[1] = res->state.n_failed; io->counts[2] = n_train;
    for (int i = 0; i < 30; i++) io->field[i] = res->field[i];
    io->field_rank = res->field_head[2];

    // ---- polyfit rounds (shopt.jl:534-577) ----
    long long nc = n_train;                          // training stars alive in the whole catalogue
    int cur = 0;
    int nan_rounds = 0;
    const bool downdate = !(getenv("SHOPT_POLY_DOWNDATE") && atoi(getenv("SHOPT_POLY_DOWNDATE")) == 0);
    bool have_raw = false;                           // acc_raw holds the previous round's sums (over nc_raw stars)
    long long nc_raw = 0;
    if (p->iterations == 0) {                        // ones(r, c, K), shopt.jl:492
        fill_kernel<<<(unsigned)(((long long)K * rc + 255) / 256), 256, 0, st>>>(io->poly, (long long)K * rc, 1.0);
        count_launches(1);
    }
    for (int loop = 1; loop <= p->iterations; loop++) {
        if (io->n_per_round) io->n_per_round[loop - 1] = nc;
        StarList sl;
        sl.idx = b.idx[cur]; sl.n = n; sl.n_dev = &b.state->nl[cur];
        if (nc < 5) {                                // interpolate.jl:142-145
            TRY(small_gather_launch(io->pgf, rc, b.idx[cur], b.pos[cur], sl.n_dev, NG, b.gbits, b.rows, ex, st));
            TRY(small_median_launch(b.rows, rc, (int)nc, p->degree, io->poly, st));
        } else {
            // Rounds after the first DOWNDATE the normal equations: [V'V | V'P] of the survivors = that of the previous
            // round minus the share of the stars it rejected (a tenth of the reads of summing the survivors again).  The
            // sums kept across rounds are the raw ones (acc_raw): a NaN there stays NaN after the subtraction, and the
            // repair below rebuilds exactly those columns from the current list, so a rejected NaN star heals itself.
            // Not used when a round dropped more than a quarter (sigma clipping can): the difference would cancel.
            const long long acc_len = (long long)K * K + (long long)K * rc;
            if (downdate && have_raw && (nc_raw - nc) * 4 <= nc_raw) {
                StarList gone;                       // this shard's share of the stars the last round dropped
                gone.idx = b.rej_idx; gone.n = std::min(n, nc_raw - nc); gone.n_dev = &b.state->n_rej;
                TRY(polyfit_normal_launch(io->pgf, rc, io->u, io->v, gone, p->degree, b.ctr, b.acc_sub, (double *)ctx->scratch,
                                          ctx->scratch_bytes, st));
                if (ex.on()) TRY(ex.sum(b.acc_sub, acc_len, 0));
                TRY(polyfit_sub_launch(b.acc_raw, b.acc_sub, acc_len, st));
            } else {
                TRY(polyfit_normal_launch(io->pgf, rc, io->u, io->v, sl, p->degree, b.ctr, b.acc_raw, (double *)ctx->scratch,
                                          ctx->scratch_bytes, st));
                if (ex.on()) TRY(ex.sum(b.acc_raw, acc_len, 0));
            }
            have_raw = true; nc_raw = nc;
            SHOPT_CUDA_TRY(cudaMemcpyAsync(b.acc, b.acc_raw, (size_t)acc_len * sizeof(double), cudaMemcpyDeviceToDevice, st));
            if (ex.on()) {
                // NaN pixel-grid values (interpolate.jl:126-130): the per-pixel median needs the values of every shard.
                // The all-reduced V'P tells every rank the same thing, so all of them enter the exchange or none does;
                // the host has to know, which costs sharded runs one stream synchronisation per round.
                if (!p->median_constraint) {
                    SHOPT_CUDA_TRY(cudaMemsetAsync(&b.state->n_nan, 0, sizeof(long long), st));
                    nan_count_kernel<<<(unsigned)((rc + 255) / 256), 256, 0, st>>>(b.acc + (size_t)K * K, rc, &b.state->n_nan);
                    count_launches(1);
                    SHOPT_CUDA_TRY(cudaMemcpyAsync(&res->n_nan, &b.state->n_nan, sizeof(long long), cudaMemcpyDeviceToHost, st));
                    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
                    if (res->n_nan > 0) {
                        TRY(polyfit_median_sharded_launch(io->pgf, rc, io->u, io->v, sl, p->degree, b.ctr, 2, b.acc, b.med,
                                                          b.shard_med, ex, st));
                        nan_rounds++;
                    }
                }
            } else if (!p->median_constraint) {      // NaN values -> median before the solve (interpolate.jl:126-130)
                TRY(polyfit_median_launch(io->pgf, rc, io->u, io->v, sl, p->degree, b.ctr, 2, b.acc, nullptr, st));
            }
            TRY(polyfit_factor_launch(b.acc, p->degree, b.ctr, nullptr, nc, 0, b.fac, st));
            if (p->median_constraint) {              // interpolate.jl:118-124, 139-141
                TRY(polyfit_factor_launch(b.acc, p->degree, b.ctr, nullptr, nc, 1, b.fac_red, st));
                if (ex.on())    // the median of z over the training stars of every shard: histogram exchange per radix pass
                    TRY(polyfit_median_sharded_launch(io->pgf, rc, io->u, io->v, sl, p->degree, b.ctr, 1, nullptr, b.med,
                                                      b.shard_med, ex, st));
                else
                    TRY(polyfit_median_launch(io->pgf, rc, io->u, io->v, sl, p->degree, b.ctr, 1, nullptr, b.med, st));
                TRY(polyfit_solve_launch(b.acc, rc, p->degree, b.fac, b.fac_red, b.med, io->poly, st));
            } else {
                TRY(polyfit_solve_launch(b.acc, rc, p->degree, b.fac, nullptr, nullptr, io->poly, st));
            }
        }
        // training errors (shopt.jl:551-555); the last round's are computed like the reference's but select nothing
        TRY(polyfit_mse_launch(io->pgf, rc, io->u, io->v, sl, p->degree, io->poly, b.mse, (double *)ctx->scratch,
                               ctx->scratch_bytes, st));
        if (loop == p->iterations) break;            // shopt.jl:557-559
        if (p->sigma_clip) {                         // shopt.jl:522-532; the count that survives is data dependent
            TRY(clip_threshold_launch(b.mse, sl.n_dev, (double)nc, b.state, b.clip, ex, st));
        } else {                                     // shopt.jl:513-520: the ceil(0.10 n) largest errors
            const long long mrej = (long long)std::ceil(0.10 * (double)nc);
            TRY(select_worst_launch(b.mse, b.pos[cur], n, sl.n_dev, mrej, b.state, b.hist, ex, st));
            nc -= mrej;
        }
        TRY(compact_training_launch(b.mse, b.pos[cur], n, cur, p->sigma_clip, b.state, b.idx[cur], b.idx[1 - cur],
                                    b.pos[1 - cur], io->flags, b.blockcnt, b.blockoff, st, b.rej_idx));
        cur = 1 - cur;
        if (p->sigma_clip) {                         // the new size of the whole catalogue's training set: one int
            long long *cnt = reinterpret_cast<long long *>(b.clip + 3);
            SHOPT_CUDA_TRY(cudaMemcpyAsync(cnt, &b.state->nl[cur], 8, cudaMemcpyDeviceToDevice, st));
            if (ex.on()) {   // int64 count as two int32 halves would carry wrongly: counts fit int32 (NG < 2^31 here)
                TRY(ex.sum(cnt, 1, 1));
            }
            SHOPT_CUDA_TRY(cudaMemcpyAsync(&res->n_after[loop - 1], cnt, 8, cudaMemcpyDeviceToHost, st));
            SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
            nc = res->n_after[loop - 1] & 0xffffffffLL;
        }
        tr.mark("  round", st);
    }
    // ---- sync 2 of 2 ----
    SHOPT_CUDA_TRY(cudaMemcpyAsync(&res->state, b.state, sizeof(StageState), cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(res->fac_head, b.fac, 16, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
    tr.mark("rounds done", st);
    io->counts[3] = nc;
    io->polyfit_rank = (p->iterations > 0 && n_train >= 5) ? res->fac_head[2] : 0;
    io->polyfit_info = (p->iterations > 0 && n_train >= 5) ? res->fac_head[3] : 0;
    if (io->polyfit_info & SHOPT_POLY_NONFINITE) set_last_error_text("polyfit: non-finite normal equations (NaN in u, v or the pixel-grid fits): coefficients are NaN");
    (void)nan_rounds;
    return SHOPT_OK;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

static int stage_prepare(const shopt_params *p, const shopt_stage_io *io, bool sharded, Ctx *ctx, StageBufs *b) {
    TRY(ctx->need_scratch(stage_scratch_bytes(p, io)));
    const StageBufs sz = stage_carve(nullptr, p, io, sharded);
    TRY(ctx->need_arena(sz.bytes));
    TRY(ctx->need_pinned(sizeof(StageResult)));
    *b = stage_carve(ctx->arena, p, io, sharded);
    return SHOPT_OK;
}
// optional per-star status outputs (device -> wherever the caller's arrays live)
static int stage_status_out(const shopt_stage_io *io, const StageBufs &b, cudaMemcpyKind kind, cudaStream_t st) {
    const size_t n = (size_t)io->n_local, nn = std::max<size_t>(n, 1);
    int32_t *dst[4] = {io->model_status, io->model_iters, io->learned_status, io->learned_iters};
    const int *src[4] = {b.tmp, b.tmp + nn, b.tmp + 3 * nn, b.tmp + 4 * nn};
    for (int i = 0; i < 4; i++)
        if (dst[i] && n) SHOPT_CUDA_TRY(cudaMemcpyAsync(dst[i], src[i], n * 4, kind, st));
    return SHOPT_OK;
}

int shopt_stage_run_dev(const shopt_params *p, shopt_stage_io *io, shopt_allreduce_fn comm, void *comm_user, void *stream) {
    TRY(stage_check_args(p, io));
    if (io->stamps_f32) return SHOPT_E_BADARG;                            // device-resident shards are float64
    Ctx *ctx; TRY(get_ctx(&ctx));
    cudaStream_t st = (cudaStream_t)stream;
    Exchange ex; ex.fn = comm; ex.user = comm_user; ex.stream = stream;
    StageBufs b;
    TRY(stage_prepare(p, io, ex.on(), ctx, &b));
    TRY(stage_begin(p, io, ex, b, st));
    TRY(stage_run_impl(p, io, ex, st, b, false));
    return stage_status_out(io, b, cudaMemcpyDeviceToDevice, st);
}

// Host buffers, one GPU.  The stamps go up in chunks on a copy stream while the per-star kernels of the chunks that
// have landed run on the compute stream, so the H2D transfer (13.7 GB at 100k x 131 x 131 in float64, half of that when
// the caller hands over the FITS-native float32 VIGNET cube, io->stamps_f32) hides behind the fits.
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
    const bool chisq = p->mode == SHOPT_MODE_CHISQ, f32 = io->stamps_f32 != 0;
    if (chisq && !io->weights) return SHOPT_E_BADARG;
    RawBuf dst, dpgf, dw, dx0p;
    TRY(ctx->need_cache(0, nn * rc * 8, &dst.p)); TRY(ctx->need_cache(1, nn * rc * 8, &dpgf.p));
    if (chisq) TRY(ctx->need_cache(2, nn * rc * 8, &dw.p));
    if (chisq && io->x0_pixels) TRY(ctx->need_cache(3, nn * rc * 8, &dx0p.p));
    // small per-star arrays of the shard: u | v | x0_model (3n) | x0_learned (3n) ; model (3n) | learned (3n) ; flags | pca_k
    Bump a2(nullptr);
    for (int pass = 0; pass < 2; pass++) {
        a2 = Bump(pass ? ctx->arena2 : nullptr);
        a2.take<double>(8 * nn); a2.take<double>(6 * nn); a2.take<int32_t>(2 * nn); a2.take<double>(K * rc);
        if (!pass) TRY(ctx->need_arena2(a2.off));
    }
    Bump a(ctx->arena2);
    double *small = a.take<double>(8 * nn), *dout = a.take<double>(6 * nn);
    int32_t *dint = a.take<int32_t>(2 * nn);
    double *dpoly = a.take<double>(K * rc);
    cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
    shopt_stage_io d = *io;
    d.stamps = dst.as<double>(); d.u = small; d.v = small + n; d.x0_model = small + 2 * n; d.x0_learned = small + 5 * n;
    d.model = dout; d.learned = dout + 3 * n; d.pgf = dpgf.as<double>();
    d.flags = dint; d.pca_k = io->pca_k ? dint + n : nullptr; d.poly = dpoly;
    d.weights = chisq ? dw.as<double>() : nullptr; d.x0_pixels = (chisq && io->x0_pixels) ? dx0p.as<double>() : nullptr;
    d.stamps_f32 = 0; d.perm_dev = nullptr;
    d.model_status = d.model_iters = d.learned_status = d.learned_iters = nullptr;
    TRY(stage_check_args(p, &d));
    Exchange ex; ex.fn = comm; ex.user = comm_user; ex.stream = st;
    StageBufs b;
    TRY(stage_prepare(p, &d, ex.on(), ctx, &b));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[0], st));
    SHOPT_CUDA_TRY(cudaStreamWaitEvent(cs, ctx->ev[0], 0));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small, io->u, n * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small + n, io->v, n * 8, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small + 2 * n, io->x0_model, n * 24, cudaMemcpyHostToDevice, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(small + 5 * n, io->x0_learned, n * 24, cudaMemcpyHostToDevice, st));
    TRY(stage_begin(p, &d, ex, b, st));
    // chunks: at least ~16 stars per resident CTA so the persistent kernels keep their tail short; at most NCHUNK_EV
    // The LAST chunk's kernels cannot hide behind an upload, so it is tapered: half, quarter, quarter of a regular chunk.
    long long nchunks = (long long)n / (16LL * 3 * sm_count());
    if (nchunks > Ctx::NCHUNK_EV - 2) nchunks = Ctx::NCHUNK_EV - 2;
    if (nchunks < 1) nchunks = 1;
    const long long per = ((long long)n + nchunks - 1) / nchunks;
    long long cuts[Ctx::NCHUNK_EV + 1];
    int ncuts = 0;
    for (long long ck = 0; ck < nchunks && ck * per < (long long)n; ck++) cuts[ncuts++] = ck * per;
    if (ncuts >= 2) {                                            // split [cuts[last], n) into 1/2, 1/4, 1/4
        const long long a0 = cuts[ncuts - 1], len = (long long)n - a0;
        if (len >= 4 * 1024) { cuts[ncuts++] = a0 + len / 2; cuts[ncuts] = a0 + len / 2 + len / 4; ncuts++; }
    }
    cuts[ncuts] = (long long)n;
    for (int ck = 0; ck < ncuts; ck++) {
        const long long first = cuts[ck], count = cuts[ck + 1] - first;
        if (count <= 0) continue;
        if (f32) {   // the raw float32 VIGNETs land in the (not yet written) pixel-grid region of the same chunk
            SHOPT_CUDA_TRY(cudaMemcpyAsync(dpgf.as<double>() + first * rc, (const float *)io->stamps + first * rc,
                                           (size_t)count * rc * 4, cudaMemcpyHostToDevice, cs));
        } else {
            SHOPT_CUDA_TRY(cudaMemcpyAsync(dst.as<double>() + first * rc, io->stamps + first * rc, (size_t)count * rc * 8,
                                           cudaMemcpyHostToDevice, cs));
        }
        if (chisq) SHOPT_CUDA_TRY(cudaMemcpyAsync(dw.as<double>() + first * rc, io->weights + first * rc,
                                                  (size_t)count * rc * 8, cudaMemcpyHostToDevice, cs));
        if (d.x0_pixels) SHOPT_CUDA_TRY(cudaMemcpyAsync(dx0p.as<double>() + first * rc, io->x0_pixels + first * rc,
                                                        (size_t)count * rc * 8, cudaMemcpyHostToDevice, cs));
        SHOPT_CUDA_TRY(cudaEventRecord(ctx->chunk_ev[ck], cs));
        SHOPT_CUDA_TRY(cudaStreamWaitEvent(st, ctx->chunk_ev[ck], 0));
        if (f32)     // nanMask (masks.jl:5-17) and, with unity_sum, ./ sum(nanToZero(.)) (dataPreprocessing.jl:342-346)
            TRY(ingest_launch(dpgf.as<double>() + first * rc, 1, io->r, io->c, count, p->unity_sum ? 1 : 0, io->r, 0,
                              dst.as<double>() + first * rc, nullptr, st));
        TRY(stage_per_star(p, &d, first, count, b.tmp, ctx, st));
    }
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[1], cs));    // last stamp on the device
    TRY(stage_run_impl(p, &d, ex, st, b, true));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[2], st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->model, d.model, n * 24, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->learned, d.learned, n * 24, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->flags, d.flags, n * 4, cudaMemcpyDeviceToHost, st));
    if (io->pca_k) SHOPT_CUDA_TRY(cudaMemcpyAsync(io->pca_k, d.pca_k, n * 4, cudaMemcpyDeviceToHost, st));
    if (io->pgf) SHOPT_CUDA_TRY(cudaMemcpyAsync(io->pgf, d.pgf, n * rc * 8, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaMemcpyAsync(io->poly, d.poly, K * rc * 8, cudaMemcpyDeviceToHost, st));
    TRY(stage_status_out(io, b, cudaMemcpyDeviceToHost, st));
    SHOPT_CUDA_TRY(cudaEventRecord(ctx->ev[3], st));
    for (int i = 0; i < 4; i++) io->counts[i] = d.counts[i];
    for (int i = 0; i < 30; i++) io->field[i] = d.field[i];
    io->field_rank = d.field_rank; io->polyfit_rank = d.polyfit_rank; io->polyfit_info = d.polyfit_info;
    return finish_timing(ctx);
}

}  // extern "C"

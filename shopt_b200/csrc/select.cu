// select.cu -- the stage's control flow between the heavy kernels, on the device: survivor counting, the training split
// along the caller's permutation (shopt.jl:472-478), worst_10_percent (shopt.jl:513-532) as a radix select, and the
// stable compaction of the training list.  Everything is integer work on O(N) small arrays, enqueue-only: the host
// learns the counts from one small pinned read-back instead of synchronising between rounds.
//
// Sharded runs exchange only: survivor bits along the permutation (N_global / 32 words + 3 counters, once), one
// 4096-bin histogram per radix pass (8 passes per applied selection), and for sigma clipping two doubles.
#include "api_ctx.h"
#include "select_core.h"
#include "stage_state.h"

namespace shopt {

constexpr int CB = 1024;   // elements per compaction block

// ---- flags + survivor counters ---------------------------------------------------------------------------------
// flags[i] = BLACKLISTED from the model s (shopt.jl:166-171) | PCA_FAILED (shopt.jl:305-309) | LEARNED_OOB
// (shopt.jl:384-386).  A learned fit that THREW has s = 0 from the reference's catch (shopt.jl:387-393), which sits
// OUTSIDE the bounds test of :384-386 (inside the try): such a star is not pushed to failedStars and stays in the
// training set, so the bounds test is skipped for it.
__global__ void flag_kernel(const double *__restrict__ s_model, const int *__restrict__ pca_status,
                            const double *__restrict__ s_learned, const int *__restrict__ learned_status, long long n,
                            double lo, double hi, int *__restrict__ flags, int *__restrict__ counters /* black, failed, surv */) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int f = 0;
    if (i < n) {
        const double sm = s_model[i];
        if (sm < lo || sm > hi) f |= SHOPT_FLAG_BLACKLISTED;
        if (pca_status[i] == SHOPT_PCA_FAILED) f |= SHOPT_FLAG_PCA_FAILED;
        const double sl = s_learned[i];
        if (learned_status[i] != SHOPT_FIT_THREW && (sl < lo || sl > hi)) f |= SHOPT_FLAG_LEARNED_OOB;
        flags[i] = f;
    }
    const int nb = __syncthreads_count(i < n && (f & SHOPT_FLAG_BLACKLISTED));
    const int nf = __syncthreads_count(i < n && !(f & SHOPT_FLAG_BLACKLISTED) && f);
    const int ns = __syncthreads_count(i < n && !f);
    if (threadIdx.x == 0) {
        if (nb) atomicAdd(counters + 0, nb);
        if (nf) atomicAdd(counters + 1, nf);
        if (ns) atomicAdd(counters + 2, ns);
    }
}

// ---- survivor bits along the permutation -------------------------------------------------------------------------
// bit i of lbits = (perm[i] is a star of this shard) && (its flags & 7) == 0; gbits receives the same words (the
// buffer that is all-reduced: disjoint bits, so the integer sum is the OR)
__global__ void surv_bits_kernel(const long long *__restrict__ perm, long long NG, long long off, long long n,
                                 const int *__restrict__ flags, unsigned *__restrict__ lbits, unsigned *__restrict__ gbits,
                                 int *__restrict__ bad_perm) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    bool s = false;
    if (i < NG) {
        const long long g = perm[i];
        if (g < 0 || g >= NG) *bad_perm = 1;
        else if (g >= off && g < off + n) s = (flags[g - off] & 7) == 0;
    }
    const unsigned w = __ballot_sync(0xffffffffu, s);
    const long long word = i >> 5;
    if ((threadIdx.x & 31) == 0 && word < ((NG + 31) >> 5)) { lbits[word] = w; if (gbits != lbits) gbits[word] = w; }
}

// inclusive block scan of one long long per thread (1024 threads): returns the inclusive prefix; *total = block sum
__device__ __forceinline__ long long block_scan_incl(long long v, long long *s_warp /* 33 */, long long *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const long long t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    __syncthreads();
    if (lane == 31) s_warp[warp] = v;
    __syncthreads();
    if (warp == 0) {
        long long w = (lane < nw) ? s_warp[lane] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += t;
        }
        s_warp[lane] = w;
    }
    __syncthreads();
    if (total) *total = s_warp[nw - 1];
    return v + (warp > 0 ? s_warp[warp - 1] : 0);
}

// n_train = round(training_ratio * n_surv) (half to even, like Julia's round(Int, .), shopt.jl:477); cutoff = the
// permutation position just after the n_train-th survivor.  One block.
__global__ void __launch_bounds__(1024) split_kernel(const unsigned *__restrict__ gbits, long long NG, double ratio,
                                                     const int *__restrict__ counters, StageState *__restrict__ st) {
    __shared__ long long s_warp[33];
    __shared__ long long s_cut;
    const long long W = (NG + 31) >> 5;
    const long long n_surv = counters[2];
    long long n_train = (long long)rint(ratio * (double)n_surv);
    if (n_train > n_surv) n_train = n_surv;
    if (n_train < 0) n_train = 0;
    if (threadIdx.x == 0) s_cut = (n_train == 0) ? 0 : NG;
    __syncthreads();
    long long run = 0;
    for (long long base = 0; base < W && n_train > 0; base += blockDim.x) {
        const long long w = base + threadIdx.x;
        const unsigned bits = (w < W) ? gbits[w] : 0u;
        long long tot;
        const long long incl = block_scan_incl((long long)__popc(bits), s_warp, &tot) + run;
        const long long excl = incl - __popc(bits);
        if (excl < n_train && n_train <= incl) {       // the n_train-th survivor lives in this word
            unsigned b = bits;
            long long need = n_train - excl;           // 1-based inside the word
            int pos = 0;
            while (need > 0) { pos = __ffs(b) - 1; b &= b - 1; need--; }
            s_cut = (w << 5) + pos + 1;
        }
        run += tot;
        __syncthreads();
        if (run >= n_train) break;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        st->n_black = counters[0]; st->n_failed = counters[1]; st->n_surv = n_surv;
        st->n_train = n_train; st->cutoff = s_cut; st->nc = n_train;
    }
}

// ---- stable compaction: count / scan / write ------------------------------------------------------------------------
struct PredSplit {     // position i of the permutation is a training star of this shard
    const unsigned *lbits; const StageState *st;
    __device__ bool operator()(long long i) const { return i < st->cutoff && ((lbits[i >> 5] >> (i & 31)) & 1u); }
};
struct PredKeep {      // entry t of the current training list survives the selection
    const double *mse; const unsigned *pos; const StageState *st; int sigma;
    __device__ bool operator()(long long t) const {
        if (sigma) return !(mse[t] > st->thr);                              // shopt.jl:522-532: err > mean + 3 std
        if (!st->sel.valid) return true;
        return sel_key(mse[t], pos[t]) < sel_prefix(st->sel);
    }
};

template <class Pred>
__global__ void __launch_bounds__(CB) compact_count_kernel(Pred pred, const long long *__restrict__ n_ptr, long long n_bound,
                                                           int *__restrict__ blockcnt) {
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long i = (long long)blockIdx.x * CB + threadIdx.x;
    const int c = __syncthreads_count(i < n && pred(i));
    if (threadIdx.x == 0) blockcnt[blockIdx.x] = c;
}
__global__ void __launch_bounds__(1024) compact_scan_kernel(const int *__restrict__ blockcnt, int nblocks,
                                                            long long *__restrict__ blockoff, long long *__restrict__ total_out,
                                                            long long *__restrict__ dropped_out,
                                                            const long long *__restrict__ n_ptr, long long n_bound) {
    __shared__ long long s_warp[33];
    long long run = 0;
    for (int base = 0; base < nblocks; base += 1024) {
        const int b = base + threadIdx.x;
        const long long c = (b < nblocks) ? blockcnt[b] : 0;
        long long tot;
        const long long incl = block_scan_incl(c, s_warp, &tot) + run;
        if (b < nblocks) blockoff[b] = incl - c;
        run += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (dropped_out) *dropped_out = (n_ptr ? *n_ptr : n_bound) - run;
        *total_out = run;     // last: total_out may alias n_ptr
    }
}
// kept entries go to out_idx/out_pos in order; mode 0: source = permutation positions (idx = perm[i] - off, pos = i,
// flags |= TRAINING); mode 1: source = current list (copy idx/pos; dropped entries get flags |= REJECTED and, when
// rej_idx is given, are listed there in order: entry i is dropped entry number i - slot)
template <class Pred>
__global__ void __launch_bounds__(CB) compact_write_kernel(Pred pred, long long n_bound, const long long *__restrict__ n_ptr,
                                                           const long long *__restrict__ blockoff, int mode,
                                                           const long long *__restrict__ perm, long long off,
                                                           const long long *__restrict__ in_idx, const unsigned *__restrict__ in_pos,
                                                           long long *__restrict__ out_idx, unsigned *__restrict__ out_pos,
                                                           int *__restrict__ flags, long long *__restrict__ rej_idx) {
    __shared__ int s_warp[32];
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long i = (long long)blockIdx.x * CB + threadIdx.x;
    const bool live = i < n;
    const bool keep = live && pred(i);
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
        int w = s_warp[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += t;
        }
        s_warp[lane] = w;     // inclusive
    }
    __syncthreads();
    const long long slot = blockoff[blockIdx.x] + (warp > 0 ? s_warp[warp - 1] : 0) + __popc(bal & ((1u << lane) - 1u));
    if (mode == 0) {
        if (keep) {
            const long long li = perm[i] - off;
            out_idx[slot] = li; out_pos[slot] = (unsigned)i;
            flags[li] |= SHOPT_FLAG_TRAINING;
        }
    } else if (live) {
        if (keep) { out_idx[slot] = in_idx[i]; out_pos[slot] = in_pos[i]; }
        else {
            const long long li = in_idx[i];
            flags[li] |= SHOPT_FLAG_REJECTED;
            if (rej_idx) rej_idx[i - slot] = li;
        }
    }
}

template <class Pred>
static int compact_launch(Pred pred, long long n_bound, const long long *n_ptr, int mode, const long long *perm, long long off,
                          const long long *in_idx, const unsigned *in_pos, long long *out_idx, unsigned *out_pos, int *flags,
                          int *blockcnt, long long *blockoff, long long *total_out, long long *dropped_out, cudaStream_t st,
                          long long *rej_idx = nullptr) {
    const int nblocks = (int)((n_bound + CB - 1) / CB);
    if (nblocks > 0) compact_count_kernel<Pred><<<nblocks, CB, 0, st>>>(pred, n_ptr, n_bound, blockcnt);
    compact_scan_kernel<<<1, 1024, 0, st>>>(blockcnt, nblocks, blockoff, total_out, dropped_out, n_ptr, n_bound);
    if (nblocks > 0)
        compact_write_kernel<Pred><<<nblocks, CB, 0, st>>>(pred, n_bound, n_ptr, blockoff, mode, perm, off, in_idx, in_pos,
                                                          out_idx, out_pos, flags, rej_idx);
    SHOPT_CUDA_TRY(cudaGetLastError());
    count_launches(nblocks > 0 ? 3 : 1);
    return SHOPT_OK;
}

// ---- radix select of the m-th largest key ------------------------------------------------------------------------------
__global__ void sel_init_kernel(StageState *__restrict__ st, long long m) {
    st->sel.prefix_hi = 0ULL; st->sel.prefix_lo = 0u; st->sel.valid = m > 0; st->sel.k = m;
}
__global__ void __launch_bounds__(256) sel_hist_kernel(const double *__restrict__ mse, const unsigned *__restrict__ pos,
                                                       const long long *__restrict__ n_ptr, const StageState *__restrict__ st,
                                                       int pass, unsigned *__restrict__ hist) {
    __shared__ unsigned sh[SEL_BINS];
    for (int b = threadIdx.x; b < SEL_BINS; b += blockDim.x) sh[b] = 0u;
    __syncthreads();
    const long long n = *n_ptr;
    const u128 prefix = sel_prefix(st->sel);
    if (st->sel.valid)
        for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
            const u128 k = sel_key(mse[t], pos[t]);
            if (sel_match(k, prefix, pass)) atomicAdd(&sh[sel_digit(k, pass)], 1u);
        }
    __syncthreads();
    for (int b = threadIdx.x; b < SEL_BINS; b += blockDim.x)
        if (sh[b]) atomicAdd(&hist[b], sh[b]);
}
// bins from the top until the cumulative count reaches k; the histogram is zeroed for the next pass.  One block.
__global__ void __launch_bounds__(1024) sel_pick_kernel(unsigned *__restrict__ hist, StageState *__restrict__ st, int pass) {
    __shared__ long long s_warp[33];
    constexpr int PER = SEL_BINS / 1024;
    long long h[PER], sum = 0;
#pragma unroll
    for (int j = 0; j < PER; j++) {                     // reversed bin index: thread t owns bins 4095 - (PER t + j)
        h[j] = hist[SEL_BINS - 1 - (PER * threadIdx.x + j)];
        sum += h[j];
    }
    long long tot;
    const long long incl = block_scan_incl(sum, s_warp, &tot);
    const long long k = st->sel.k;
    const bool valid = st->sel.valid;
    __syncthreads();
    if (valid && tot < k && threadIdx.x == 0) st->sel.valid = 0;
    if (valid && incl - sum < k && k <= incl) {
        long long cum = incl - sum;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            if (cum + h[j] >= k) {
                const unsigned b = (unsigned)(SEL_BINS - 1 - (PER * threadIdx.x + j));
                SelState s = st->sel;
                sel_set_prefix(s, sel_prefix(s) | ((u128)b << (96 - SEL_BITS * (pass + 1))));
                s.k = k - cum;
                st->sel = s;
                break;
            }
            cum += h[j];
        }
    }
#pragma unroll
    for (int j = 0; j < PER; j++) hist[SEL_BINS - 1 - (PER * threadIdx.x + j)] = 0u;
}

// ---- sigma clipping: err > mean + 3 std (sample std), shopt.jl:522-532 ------------------------------------------------
// one block, fixed-order sums: out[0] = sum_t f(mse[t]) with f = identity (phase 0) or (e - mean)^2 (phase 1)
__global__ void __launch_bounds__(1024) clip_sum_kernel(const double *__restrict__ mse, const long long *__restrict__ n_ptr,
                                                        int phase, double nc, double *__restrict__ out) {
    __shared__ double s_part[32];
    const long long n = *n_ptr;
    const double mean = phase ? out[0] / nc : 0.0;
    double a = 0.0;
    for (long long t = threadIdx.x; t < n; t += 1024) { const double e = mse[t] - mean; a += phase ? e * e : e; }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) a += __shfl_xor_sync(0xffffffffu, a, m);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 32; w++) s += s_part[w];
        out[1 + phase] = s;
    }
}
__global__ void clip_thr_kernel(const double *__restrict__ sums, double nc, StageState *__restrict__ st) {
    const double mean = sums[1] / nc;
    st->thr = mean + 3.0 * sqrt(sums[2] / (nc - 1.0));
}
__global__ void clip_mean_kernel(double *__restrict__ sums) { sums[0] = sums[1]; }   // [0] = all-reduced sum of errors

// ---- (u, v) centre and scale for the interpolation basis ---------------------------------------------------------------
// sums[0..5) = n, sum u, sum v, sum u^2, sum v^2 over the shard (fixed order); all-reduced; centre = mean, scale such
// that a uniform distribution maps to [-1, 1].  Any affine map gives the same polynomial; this one keeps V'V
// well-conditioned when the caller passes raw sky coordinates (RA ~ 150 deg, Dec ~ 2 deg, field 0.2 deg).
__global__ void __launch_bounds__(1024) uv_sums_kernel(const double *__restrict__ u, const double *__restrict__ v, long long n,
                                                       double *__restrict__ sums) {
    __shared__ double s_part[32][4];
    double a[4] = {0, 0, 0, 0};
    for (long long t = threadIdx.x; t < n; t += 1024) {
        const double x = u[t], y = v[t];
        a[0] += x; a[1] += y; a[2] += x * x; a[3] += y * y;
    }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1)
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] += __shfl_xor_sync(0xffffffffu, a[i], m);
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int i = 0; i < 4; i++) s_part[threadIdx.x >> 5][i] = a[i];
    __syncthreads();
    if (threadIdx.x < 4) {
        double s = 0.0;
        for (int w = 0; w < 32; w++) s += s_part[w][threadIdx.x];
        sums[1 + threadIdx.x] = s;
    }
    if (threadIdx.x == 0) sums[0] = (double)n;
}
__global__ void uv_centre_kernel(const double *__restrict__ sums, double *__restrict__ ctr /* u0, 1/su, v0, 1/sv */) {
    const double n = sums[0];
    double u0 = 0.0, v0 = 0.0, su = 1.0, sv = 1.0;
    if (n > 0.0) {
        u0 = sums[1] / n; v0 = sums[2] / n;
        const double vu = sums[3] / n - u0 * u0, vv = sums[4] / n - v0 * v0;
        if (isfinite(vu) && vu > 1e-12 * u0 * u0 && vu > 0.0) su = sqrt(3.0 * vu);
        if (isfinite(vv) && vv > 1e-12 * v0 * v0 && vv > 0.0) sv = sqrt(3.0 * vv);
    }
    if (!isfinite(u0)) u0 = 0.0;
    if (!isfinite(v0)) v0 = 0.0;
    ctr[0] = u0; ctr[1] = 1.0 / su; ctr[2] = v0; ctr[3] = 1.0 / sv;
}

// ---- launches ---------------------------------------------------------------------------------------------------------
int flag_launch(const double *s_model, const int *pca_status, const double *s_learned, const int *learned_status, long long n,
                double lo, double hi, int *flags, int *counters, cudaStream_t st) {
    SHOPT_CUDA_TRY(cudaMemsetAsync(counters, 0, 3 * sizeof(int), st));
    if (n > 0) {
        flag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s_model, pca_status, s_learned, learned_status, n, lo, hi,
                                                                 flags, counters);
        count_launches(1);
    }
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int surv_bits_launch(const long long *perm, long long NG, long long off, long long n, const int *flags, unsigned *lbits,
                     unsigned *gbits, int *bad_perm, cudaStream_t st) {
    if (NG > 0) {
        surv_bits_kernel<<<(unsigned)((NG + 255) / 256), 256, 0, st>>>(perm, NG, off, n, flags, lbits, gbits, bad_perm);
        count_launches(1);
    }
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int split_launch(const unsigned *gbits, long long NG, double ratio, const int *counters, StageState *state, cudaStream_t st) {
    split_kernel<<<1, 1024, 0, st>>>(gbits, NG, ratio, counters, state);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int training_list_launch(const unsigned *lbits, const long long *perm, long long NG, long long off, StageState *state,
                         long long *out_idx, unsigned *out_pos, int *flags, int *blockcnt, long long *blockoff,
                         cudaStream_t st) {
    PredSplit pred{lbits, state};
    return compact_launch(pred, NG, nullptr, 0, perm, off, nullptr, nullptr, out_idx, out_pos, flags, blockcnt, blockoff,
                          &state->nl[0], nullptr, st);
}
int select_worst_launch(const double *mse, const unsigned *pos, long long n_bound, const long long *n_ptr, long long m,
                        StageState *state, unsigned *hist, const Exchange &ex, cudaStream_t st) {
    sel_init_kernel<<<1, 1, 0, st>>>(state, m);
    count_launches(1);
    if (m <= 0) { SHOPT_CUDA_TRY(cudaGetLastError()); return SHOPT_OK; }
    int grid = (int)((n_bound + 256 * 8 - 1) / (256 * 8));
    if (grid < 1) grid = 1;
    if (grid > 2 * sm_count()) grid = 2 * sm_count();
    for (int pass = 0; pass < SEL_PASSES; pass++) {
        sel_hist_kernel<<<grid, 256, 0, st>>>(mse, pos, n_ptr, state, pass, hist);
        TRY(ex.sum(hist, SEL_BINS, 1));
        sel_pick_kernel<<<1, 1024, 0, st>>>(hist, state, pass);
        count_launches(2);
    }
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int clip_threshold_launch(const double *mse, const long long *n_ptr, double nc, StageState *state, double *sums /* 3 */,
                          const Exchange &ex, cudaStream_t st) {
    clip_sum_kernel<<<1, 1024, 0, st>>>(mse, n_ptr, 0, nc, sums);
    TRY(ex.sum(sums + 1, 1, 0));
    clip_mean_kernel<<<1, 1, 0, st>>>(sums);
    clip_sum_kernel<<<1, 1024, 0, st>>>(mse, n_ptr, 1, nc, sums);
    TRY(ex.sum(sums + 2, 1, 0));
    clip_thr_kernel<<<1, 1, 0, st>>>(sums, nc, state);
    count_launches(4);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int compact_training_launch(const double *mse, const unsigned *pos, long long n_bound, int cur, int sigma, StageState *state,
                            const long long *in_idx, long long *out_idx, unsigned *out_pos, int *flags, int *blockcnt,
                            long long *blockoff, cudaStream_t st, long long *rej_idx) {
    PredKeep pred{mse, pos, state, sigma};
    return compact_launch(pred, n_bound, &state->nl[cur], 1, nullptr, 0, in_idx, pos, out_idx, out_pos, flags, blockcnt,
                          blockoff, &state->nl[1 - cur], &state->n_rej, st, rej_idx);
}
int uv_centre_launch(const double *u, const double *v, long long n, double *sums, double *ctr, const Exchange &ex,
                     cudaStream_t st) {
    uv_sums_kernel<<<1, 1024, 0, st>>>(u, v, n, sums);
    TRY(ex.sum(sums, 5, 0));
    uv_centre_kernel<<<1, 1, 0, st>>>(sums, ctr);
    count_launches(2);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

}  // namespace shopt

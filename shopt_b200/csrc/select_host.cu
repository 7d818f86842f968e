// select_host.cu -- HOST restatement of the sharded stage's exchange protocol (select.cu), for hosts without a device:
// the world_size-2 gloo tests drive it through a host all-reduce callback and check that every rank derives exactly
// the single-process training split (shopt.jl:472-478) and rejection sequence (shopt.jl:513-520, 557-575), including
// the stable tie order.  Same order relations (select_core.h), same messages as the device path:
//   split:   N_global / 32 words of survivor bits along the permutation + 3 counters, one sum-all-reduce;
//   select:  8 radix passes, each one sum-all-reduce of a 4096-bin histogram;
//   median:  per pixel, 2 x 8 radix passes of 256 bins (lower and upper middle order statistic; polyfit.cu shard_med_*).
#include <cmath>
#include <cstring>
#include <vector>

#include "api_ctx.h"
#include "select_core.h"

using namespace shopt;

extern "C" {

// flags_local (n_local): SHOPT_FLAG_* of this shard's stars [offset, offset + n_local); perm (n_global).
// train_idx_out / train_pos_out (capacity n_local): this shard's training stars (LOCAL indices) in training order and
// their positions in the permutation.  counts_out[4] = blacklisted, failed, survivors, training stars of the whole
// catalogue.  Returns the number of local training stars or a negative SHOPT_E_*.
int64_t shopt_training_split_sharded_host(const int32_t *flags_local, int64_t n_local, int64_t offset, int64_t n_global,
                                          const int64_t *perm, double training_ratio, shopt_host_allreduce_fn allreduce,
                                          void *user, int64_t *train_idx_out, uint32_t *train_pos_out, int64_t *counts_out) {
    if (n_local < 0 || n_global < n_local || offset < 0 || offset + n_local > n_global || (n_global > 0 && !perm) ||
        (n_local > 0 && (!flags_local || !train_idx_out || !train_pos_out)))
        return SHOPT_E_BADARG;
    const int64_t W = (n_global + 31) >> 5;
    std::vector<uint32_t> lbits((size_t)W, 0u), gbits((size_t)W + 3, 0u);
    int32_t *counters = reinterpret_cast<int32_t *>(gbits.data() + W);
    for (int64_t i = 0; i < n_local; i++) {                       // flag_kernel's counters
        const int f = flags_local[i] & 7;
        if (f & SHOPT_FLAG_BLACKLISTED) counters[0]++;
        else if (f) counters[1]++;
        else counters[2]++;
    }
    for (int64_t i = 0; i < n_global; i++) {                      // surv_bits_kernel
        const int64_t g = perm[i];
        if (g < 0 || g >= n_global) return SHOPT_E_BADARG;
        if (g >= offset && g < offset + n_local && (flags_local[g - offset] & 7) == 0) lbits[(size_t)(i >> 5)] |= 1u << (i & 31);
    }
    std::memcpy(gbits.data(), lbits.data(), (size_t)W * 4);
    if (allreduce && allreduce(gbits.data(), W + 3, 1, user) != 0) return SHOPT_E_CUDA;
    const int64_t n_surv = counters[2];                           // split_kernel
    int64_t n_train = (int64_t)std::nearbyint(training_ratio * (double)n_surv);
    if (n_train > n_surv) n_train = n_surv;
    if (n_train < 0) n_train = 0;
    int64_t cutoff = (n_train == 0) ? 0 : n_global, run = 0;
    for (int64_t w = 0; w < W && n_train > 0; w++) {
        const uint32_t bits = gbits[(size_t)w];
        const int pc = __builtin_popcount(bits);
        if (run < n_train && n_train <= run + pc) {
            uint32_t b = bits;
            int pos = 0;
            for (int64_t need = n_train - run; need > 0; need--) { pos = __builtin_ctz(b); b &= b - 1; }
            cutoff = (w << 5) + pos + 1;
            break;
        }
        run += pc;
    }
    int64_t m = 0;                                                // training_list (compaction along the permutation)
    for (int64_t i = 0; i < cutoff; i++)
        if ((lbits[(size_t)(i >> 5)] >> (i & 31)) & 1u) { train_idx_out[m] = perm[i] - offset; train_pos_out[m] = (uint32_t)i; m++; }
    if (counts_out) { counts_out[0] = counters[0]; counts_out[1] = counters[1]; counts_out[2] = n_surv; counts_out[3] = n_train; }
    return m;
}

// err_local / pos_local (n_local): this shard's training errors and permutation positions; m = ceil(0.10 n) of the WHOLE
// training set.  reject_out[t] = 1 when entry t is among the m worst of the whole catalogue (error descending, NaN
// first, ties by position).  Returns the number of local rejections or a negative SHOPT_E_*.
int64_t shopt_select_worst_sharded_host(const double *err_local, const uint32_t *pos_local, int64_t n_local, int64_t m,
                                        shopt_host_allreduce_fn allreduce, void *user, uint8_t *reject_out) {
    if (n_local < 0 || m < 0 || (n_local > 0 && (!err_local || !pos_local || !reject_out))) return SHOPT_E_BADARG;
    SelState st;
    st.prefix_hi = 0; st.prefix_lo = 0; st.valid = m > 0; st.k = m;
    std::vector<uint32_t> hist(SEL_BINS);
    for (int pass = 0; pass < SEL_PASSES && m > 0; pass++) {
        std::fill(hist.begin(), hist.end(), 0u);
        const u128 prefix = sel_prefix(st);
        if (st.valid)
            for (int64_t t = 0; t < n_local; t++) {
                const u128 k = sel_key(err_local[t], pos_local[t]);
                if (sel_match(k, prefix, pass)) hist[sel_digit(k, pass)]++;
            }
        if (allreduce && allreduce(hist.data(), SEL_BINS, 1, user) != 0) return SHOPT_E_CUDA;
        if (st.valid) sel_pick_host(hist.data(), st, pass);
    }
    int64_t nrej = 0;
    for (int64_t t = 0; t < n_local; t++) {
        const bool keep = !st.valid || sel_key(err_local[t], pos_local[t]) < sel_prefix(st);
        reject_out[t] = keep ? 0 : 1;
        nrej += !keep;
    }
    return nrej;
}

// median(filter(!isnan, z)) of interpolate.jl:119,128,140 per pixel when the training stars are sharded: z_local is this
// rank's (n_local, npx) block, row-major (star, pixel).  Every rank returns the same med_out (npx): NaN for a pixel
// without finite values, Statistics.middle(v1, v2) = v1/2 + v2/2 for an even count.  The exchange is one int32
// sum-all-reduce of npx x 256 histogram bins per pass, 16 passes: what polyfit_median_sharded_launch does on the device.
int shopt_median_sharded_host(const double *z_local, int64_t n_local, int64_t npx, shopt_host_allreduce_fn allreduce, void *user,
                              double *med_out) {
    if (n_local < 0 || npx < 0 || (npx > 0 && !med_out) || (n_local > 0 && npx > 0 && !z_local)) return SHOPT_E_BADARG;
    auto order_key = [](double z) {
        unsigned long long b; std::memcpy(&b, &z, 8);
        return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
    };
    auto key_value = [](unsigned long long k) {
        const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffULL) : ~k;
        double z; std::memcpy(&z, &b, 8); return z;
    };
    std::vector<unsigned long long> prefix((size_t)2 * npx, 0ULL);
    std::vector<uint32_t> kleft((size_t)2 * npx, 0u), total((size_t)npx, 0u), hist((size_t)npx * 256);
    for (int which = 0; which < 2; which++)
        for (int pass = 0; pass < 8; pass++) {
            const int shift = 56 - 8 * pass;
            std::fill(hist.begin(), hist.end(), 0u);
            for (int64_t t = 0; t < n_local; t++)
                for (int64_t px = 0; px < npx; px++) {
                    const double z = z_local[(size_t)t * npx + px];
                    if (std::isnan(z)) continue;
                    const unsigned long long key = order_key(z), pre = prefix[(size_t)which * npx + px];
                    if (pass == 0 || (key >> (shift + 8)) == (pre >> (shift + 8))) hist[(size_t)px * 256 + ((key >> shift) & 255ULL)]++;
                }
            if (allreduce && npx > 0 && allreduce(hist.data(), npx * 256, 1, user) != 0) return SHOPT_E_CUDA;
            for (int64_t px = 0; px < npx; px++) {
                const uint32_t *h = &hist[(size_t)px * 256];
                uint32_t k = kleft[(size_t)which * npx + px];
                if (pass == 0) {
                    uint32_t tot = 0;
                    for (int b = 0; b < 256; b++) tot += h[b];
                    total[(size_t)px] = tot;
                    k = tot ? (which ? tot / 2 : (tot - 1) / 2) : 0;
                }
                uint32_t cum = 0;
                for (int b = 0; b < 256; b++) {
                    if (cum + h[b] > k) { prefix[(size_t)which * npx + px] |= (unsigned long long)b << shift; k -= cum; break; }
                    cum += h[b];
                }
                kleft[(size_t)which * npx + px] = k;
            }
        }
    for (int64_t px = 0; px < npx; px++) {
        if (!total[(size_t)px]) { med_out[px] = std::nan(""); continue; }
        const double v1 = key_value(prefix[(size_t)px]), v2 = key_value(prefix[(size_t)npx + px]);
        med_out[px] = (total[(size_t)px] & 1u) ? v1 : v1 / 2 + v2 / 2;
    }
    return SHOPT_OK;
}

}  // extern "C"

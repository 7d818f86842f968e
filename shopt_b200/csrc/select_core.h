// select_core.h -- the order relations of the stage's selections, shared by the device kernels (select.cu) and the
// host restatement of the sharded protocol (shopt_select_worst_sharded_host, driven by the gloo tests):
//   worst_10_percent (shopt.jl:513-520): sort (index, error) by error, rev = true, stable -> error descending with NaN
//   first (isless treats NaN as the largest), ties keep the lower index (= earlier position in the training order)
//   first; the first ceil(0.10 n) are rejected.
// The total order is turned into ONE unsigned 96-bit key (error key : ~position); since positions are unique the keys
// are, and the rejected set is { key >= T } with T the m-th largest key, found by an MSB-first radix select whose
// per-pass histogram (4096 bins) is the only thing ranks exchange.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define SHOPT_HD __host__ __device__ __forceinline__
#else
#define SHOPT_HD inline
#endif

namespace shopt {

typedef unsigned __int128 u128;
constexpr int SEL_BITS = 12, SEL_BINS = 1 << SEL_BITS, SEL_PASSES = 8;   // 8 x 12 = 96 bits

SHOPT_HD unsigned long long f64_bits(double z) {
#if defined(__CUDA_ARCH__)
    return (unsigned long long)__double_as_longlong(z);
#else
    union { double d; unsigned long long u; } c; c.d = z; return c.u;
#endif
}
// monotone map of the doubles under isless (-Inf < ... < -0.0 < 0.0 < ... < Inf < NaN) onto unsigned integers
SHOPT_HD unsigned long long err_key(double e) {
    if (e != e) return ~0ULL;
    const unsigned long long b = f64_bits(e);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
SHOPT_HD u128 sel_key(double e, unsigned pos) { return ((u128)err_key(e) << 32) | (u128)(unsigned)(~pos); }
SHOPT_HD unsigned sel_digit(u128 k, int pass) { return (unsigned)(k >> (96 - SEL_BITS * (pass + 1))) & (SEL_BINS - 1); }
SHOPT_HD bool sel_match(u128 k, u128 prefix, int pass) {
    return pass == 0 || (k >> (96 - SEL_BITS * pass)) == (prefix >> (96 - SEL_BITS * pass));
}

// state of one selection: after SEL_PASSES passes `prefix` is the m-th largest key; k = rank still to find inside the
// matching keys (1-based); valid = 0 when m == 0 (nothing is rejected)
struct SelState {
    unsigned long long prefix_hi;   // bits 95..32
    unsigned int prefix_lo;         // bits 31..0
    int valid;
    long long k;
};
SHOPT_HD u128 sel_prefix(const SelState &s) { return ((u128)s.prefix_hi << 32) | (u128)s.prefix_lo; }
SHOPT_HD void sel_set_prefix(SelState &s, u128 p) { s.prefix_hi = (unsigned long long)(p >> 32); s.prefix_lo = (unsigned)p; }

// sequential form of the pick (host restatement; the device kernel does the same with a block scan): bins from the top
inline void sel_pick_host(const unsigned *hist, SelState &s, int pass) {
    long long cum = 0;
    for (int b = SEL_BINS - 1; b >= 0; b--) {
        const long long h = hist[b];
        if (cum + h >= s.k) {
            sel_set_prefix(s, sel_prefix(s) | ((u128)(unsigned)b << (96 - SEL_BITS * (pass + 1))));
            s.k -= cum;
            return;
        }
        cum += h;
    }
    s.valid = 0;   // fewer than k keys: cannot happen when m <= n
}

}  // namespace shopt

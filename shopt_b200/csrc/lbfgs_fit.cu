// lbfgs_fit.cu -- batched analytic profile fit: one persistent CTA per star.
//
// Replaces the loops shopt/shopt.jl:130-164 and :357-397 of the reference:
//     Optim.optimize(cost, g!, x0, LBFGS(), Optim.Options(g_tol = ...))      (shopt.jl:139-143, 367-371)
// with cost/costD = shopt/analyticLBFGS.jl:8-46 / 49-87 (sum-normalised elliptical Gaussian, fGaussian of
// shopt/radialProfiles.jl:5-9, on the get_middle_nxn crop dataPreprocessing.jl:207-216) and
// g!/gD! = ForwardDiff.gradient (analyticLBFGS.jl:89-101).  Optim 1.7.8 / LineSearches 7.2.0 are not
// vendored in the reference; the algorithm followed is the one recorded in SURVEY.md Appendix A.
//
// Design (B200): persistent CTAs pull stars from a device counter; the n_a x n_a crop of a stamp is staged into shared
// memory by 1-D bulk (TMA) copies, one per crop column.  A CTA serves NSLOT stars at a time: each star has a MASTER
// warp that owns the scalar LBFGS / HagerZhang state machine (history and line-search points in shared-memory slots)
// while the CTA's WORKER warps alternate between the stars on named barriers and run the two pixel passes of whichever
// value+gradient evaluation is requested (profile moments over all pixels, residual moments over the non-NaN ones; the
// per-pixel profile value is parked in a shared-memory plane between the passes; exp is table-driven and branch-free).
// One star's serial phase hides behind the other star's pixel passes.  Reductions are deterministic (fixed halving
// trees, warp partials summed in warp order).  The manifold reparameterisation [sigma,e1,e2] -> [s,g1,g2] is evaluated
// with explicitly rounded operations (no FMA contraction) in the same operation order as the reference's expression,
// because that expression amplifies round-off by 1/|e|^4; the dual-number partials follow ForwardDiff
// op for op, regrouped over the master warp's lanes without changing a single operation.
// Measured and dropped: pass A by the geometric recurrence f(v + TV) = f(v) r(v), r(v + TV) = r(v) w outwards from the
// parabola's vertex (two multiplications per pixel instead of a 12-operation exp, three seed exps per thread): 42.1 ms
// against 41.9 ms per 100k stars and the 1e-12 cost parity lost -- the kernel is bound by dependent-operation latency in
// the pixel passes and by the master / worker hand-offs, not by its FP64 operation count.
// Also measured and dropped: the pairs of the symmetric pass A with their profile values kept in REGISTERS across the
// barrier and pass B walking the same pairs (no plane stores / loads, one set of moment updates per pair: ~30 % fewer
// instructions in the pixel passes on paper): 49.2 ms against 37.7 ms per 100k stars -- eight live doubles per thread
// across the barrier and the fully unrolled guarded pair loop cost more than the plane traffic they replace, and merely
// compiling that path in slowed the plane path to 42.8 ms (register allocation of the shared worker loop).
#include <cstdlib>
#include <type_traits>
#include "common.cuh"

namespace shopt {

// ---------------------------------------------------------------------------------------------
// Scalar section: dual numbers with 3 partials, evaluated op for op like ForwardDiff does on the
// reference's expression.  __dmul_rn/__dadd_rn/... are never contracted into FMAs.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double mul_(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double div_(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double sqrt_(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ double bcast_(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
// Float64 ^ 4 the way Julia >= 1.8 evaluates a literal integer power (compensated squaring).
__device__ __forceinline__ double pow4_(double x) {
    double x2 = mul_(x, x);
    double lo2 = __fma_rn(x, x, -x2);
    double err = mul_(mul_(x2, 2.0), lo2);
    double x4 = mul_(x2, x2);
    double lo4 = add_(__fma_rn(x2, x2, -x4), err);
    if (!isfinite(x4)) return x4;
    return add_(x4, lo4);
}
// ---------------------------------------------------------------------------------------------
// analyticLBFGS.jl:10-18 under ForwardDiff, restructured for ONE warp without changing a single IEEE operation.
// ForwardDiff's Dual{3} arithmetic on  sigma, e1, e2 -> g1, g2, ms = s^2 / (1 - g1^2 - g2^2)  is 24 IEEE divisions and
// square roots in a dependent-looking chain; executed by every lane redundantly (round 1) it WAS the critical path of
// an evaluation: the worker warps sat on the barrier while the master ground through it.  But
//   * the VALUE channel needs only 10 of them, three of which are independent (one lane-parallel division), as are
//     g1 and g2 (another);
//   * every other division is a helper of a partial (1/b, a/b^2, 1/(2 sqrt), (a/b)/b) whose operands are VALUES: all
//     thirteen are independent of each other and run as ONE division, one helper per lane;
//   * the three partial channels are independent: lane c carries channel c, and what is left per channel is a
//     straight line of multiplications and additions.
// Same operations on the same operands with the same roundings (mul_/add_/... are never contracted): the bits are
// those of the op-for-op dual evaluation.
// ---------------------------------------------------------------------------------------------
struct ScalarVals {   // value channel with the intermediates the partials need
    double ell, ell2, p4, r4, r1p, r1, inner, normG, ratio, g1, g2, sg, sg2, D, ms;
};
__device__ __forceinline__ ScalarVals scalar_values(double x0, double x1, double x2, int lane) {
    ScalarVals v;
    v.ell = sqrt_(add_(mul_(x1, x1), mul_(x2, x2)));
    v.ell2 = mul_(v.ell, v.ell);
    v.p4 = pow4_(v.ell);
    {   // 4 / ell2 (lane 0), 1 / ell^4 (lane 1), 1 / ell2 (lane 2)
        const double q = div_((lane == 0) ? 4.0 : 1.0, (lane == 1) ? v.p4 : v.ell2);
        v.r4 = bcast_(q, 0); v.r1p = bcast_(q, 1); v.r1 = bcast_(q, 2);
    }
    v.inner = sqrt_(add_(v.r4, v.r1p));
    v.normG = sqrt_(add_(1.0, mul_(0.5, sub_(v.r1, v.inner))));
    v.ratio = div_(v.ell, v.normG);
    {   // g1 = e1 / ratio (lane 0), g2 = e2 / ratio (lane 1)
        const double q = div_((lane == 0) ? x1 : x2, v.ratio);
        v.g1 = bcast_(q, 0); v.g2 = bcast_(q, 1);
    }
    v.sg = mul_(x0, x0);
    v.sg2 = mul_(v.sg, v.sg);
    v.D = sub_(sub_(1.0, mul_(v.g1, v.g1)), mul_(v.g2, v.g2));
    v.ms = div_(v.sg2, v.D);
    return v;
}
// partial channel `lane` (0: d/dsigma, 1: d/de1, 2: d/de2; other lanes carry zeros) of a11 = 1 + g1, g2 and ms
// (a22 = 1 - g1 has the negated partial of a11)
__device__ __forceinline__ void scalar_partials(const ScalarVals &v, double x0, double x1, double x2, int lane, double &a11_d,
                                                double &g2_d, double &ms_d) {
    double num, den;
    switch (lane) {   // the helper quotient of lane k
        case 0: num = 1.0; den = mul_(2.0, v.ell); break;                  // d_sqrt(sum):        1 / (2 ell)
        case 1: num = v.r4; den = v.ell2; break;                           // d_rdiv(4, ell2):   -(4/ell2) / ell2
        case 2: num = v.r1p; den = v.p4; break;                            // d_rdiv(1, ell^4):  -(1/ell^4) / ell^4
        case 3: num = 1.0; den = mul_(2.0, v.inner); break;                // d_sqrt(...):        1 / (2 inner)
        case 4: num = v.r1; den = v.ell2; break;                           // d_rdiv(1, ell2):   -(1/ell2) / ell2
        case 5: num = 1.0; den = mul_(2.0, v.normG); break;                // d_sqrt(...):        1 / (2 normG)
        case 6: num = 1.0; den = v.normG; break;                           // d_div(ell, normG):  1 / normG
        case 7: num = v.ell; den = mul_(v.normG, v.normG); break;          //                    -ell / normG^2
        case 8: num = 1.0; den = v.ratio; break;                           // d_div(e, ratio):    1 / ratio
        case 9: num = x1; den = mul_(v.ratio, v.ratio); break;             //                    -e1 / ratio^2
        case 10: num = x2; den = mul_(v.ratio, v.ratio); break;            //                    -e2 / ratio^2
        case 11: num = 1.0; den = v.D; break;                              // d_div(sg^2, D):     1 / D
        case 12: num = v.sg2; den = mul_(v.D, v.D); break;                 //                    -sg^2 / D^2
        default: num = 1.0; den = 1.0; break;
    }
    const double q = div_(num, den);
    const double w_ell = bcast_(q, 0), w4 = -bcast_(q, 1), w1p = -bcast_(q, 2), w_inner = bcast_(q, 3), w1 = -bcast_(q, 4);
    const double w_ng = bcast_(q, 5), ib_ng = bcast_(q, 6), w_ratio = -bcast_(q, 7), ib_r = bcast_(q, 8);
    const double w_g1 = -bcast_(q, 9), w_g2 = -bcast_(q, 10), ib_D = bcast_(q, 11), w_ms = -bcast_(q, 12);
    const double sig_d = (lane == 0) ? 1.0 : 0.0, e1_d = (lane == 1) ? 1.0 : 0.0, e2_d = (lane == 2) ? 1.0 : 0.0;
    const double sg_d = mul_(mul_(2.0, x0), sig_d);                                  // d_sq(sig)
    const double sum_d = add_(mul_(mul_(2.0, x1), e1_d), mul_(mul_(2.0, x2), e2_d)); // d_add(d_sq(e1), d_sq(e2))
    const double ell_d = mul_(w_ell, sum_d);                                         // d_sqrt
    const double ell2_d = mul_(mul_(2.0, v.ell), ell_d);                             // d_sq(ell)
    const double r4_d = mul_(w4, ell2_d);                                            // d_rdiv(4, ell2)
    const double p4_d = mul_(mul_(ell_d, 4.0), mul_(mul_(v.ell, v.ell), v.ell));     // d_pow4(ell)
    const double r1p_d = mul_(w1p, p4_d);                                            // d_rdiv(1, ell^4)
    const double inner_d = mul_(w_inner, add_(r4_d, r1p_d));                         // d_sqrt(d_add)
    const double r1_d = mul_(w1, ell2_d);                                            // d_rdiv(1, ell2)
    const double ng2_d = mul_(sub_(r1_d, inner_d), 0.5);                             // d_radd(1, d_rmul(0.5, d_sub))
    const double normG_d = mul_(w_ng, ng2_d);                                        // d_sqrt
    const double ratio_d = add_(mul_(ell_d, ib_ng), mul_(normG_d, w_ratio));         // d_div(ell, normG)
    const double g1_d = add_(mul_(e1_d, ib_r), mul_(ratio_d, w_g1));                 // d_div(e1, ratio)
    g2_d = add_(mul_(e2_d, ib_r), mul_(ratio_d, w_g2));                              // d_div(e2, ratio)
    a11_d = g1_d;                                                                    // d_radd(1, g1); a22: d_rsub -> -g1_d
    const double sg2_d = mul_(mul_(2.0, v.sg), sg_d);                                // d_sq(sg)
    const double D_d = sub_(-mul_(mul_(2.0, v.g1), g1_d), mul_(mul_(2.0, v.g2), g2_d));   // d_sub(d_rsub(1, d_sq(g1)), d_sq(g2))
    ms_d = add_(mul_(sg2_d, ib_D), mul_(D_d, w_ms));                                 // d_div(sg^2, D)
}
// shopt.jl:145-153 on plain doubles
__device__ __forceinline__ void e_to_g(double e1, double e2, double &g1, double &g2) {
    double ell = sqrt_(add_(mul_(e1, e1), mul_(e2, e2)));
    double ell2 = mul_(ell, ell);
    double normG = sqrt_(add_(1.0, mul_(0.5, sub_(div_(1.0, ell2), sqrt_(add_(div_(4.0, ell2), div_(1.0, pow4_(ell))))))));
    double ratio = div_(ell, normG);
    g1 = div_(e1, ratio);
    g2 = div_(e2, ratio);
}

// ---------------------------------------------------------------------------------------------
// One CTA's view of the star being fitted.
//
// Thread layout: 2-D, thread (tu, tv) owns row u = tu of the crop and walks the columns v = tv, tv + TV, ... so that
// du is a per-thread constant and dv advances by a constant -- no integer division or int->double conversion per
// pixel.  With du fixed, the exponent  kappa q(du,dv) = c0 + dv (c1 + cc dv)  is two FMAs per pixel, and every sum
// the value and gradient need is a MOMENT of the profile: q, w1 du, w1 dv + w2 du, w2 dv are all linear
// combinations of du^2, du dv, dv^2, so the passes accumulate only
//     pass A (all pixels):      sum f, sum f dv, sum f dv^2                 (x du^2, du applied once per thread)
//     pass B (non-NaN pixels):  sum rho^2, sum rho f, sum rho f dv, sum rho f dv^2
// and the scalar chain rule rebuilds the reference's sums from them.  The per-pixel profile value is parked in a
// shared-memory plane between the passes (registers stay free; three CTAs fit an SM).
// The dual-number scalar section (the manifold reparameterisation, ~10 IEEE divisions/square roots on 4-vectors)
// is evaluated by warp 0 only and broadcast through shared memory.
// ---------------------------------------------------------------------------------------------
// exp(x) for the profile, x = -kappa q <= 0, table-driven: 12 FP64 operations against 19 for a table-free Taylor form
// and ~35 for the library's (the FP64 pipe is what bounds this kernel: profiles/).  n = rint(x 64 / ln 2) by the
// 1.5 * 2^52 shift, r = x - n ln2/64 in two pieces (|r| <= 0.00542), exp(x) = 2^(n >> 6) T[n & 63] (1 + r + r^2 P(r))
// with T[i] = 2^(i/64) correctly rounded and P the degree-3 interpolant of (e^r - 1 - r) / r^2 at Chebyshev nodes
// (relative error of the polynomial 4.4e-18).  Scaling by a multiplication with 2^e built in the exponent field, so NaN
// propagates; results below the normal range are flushed to 0.  Within one ulp, like the library's.  Branch-free, so
// that the pixels a thread has in flight interleave.  Valid for x <= 709 (the callers route evaluations whose kappa is
// negative, infinite or NaN -- never produced by a finite profile -- to the library's exp).
// The table lives in shared memory, replicated 16 times so that lane l reads copy l & 15 -- bank l & 15 of 16 eight-byte
// banks: two wavefronts per warp whatever the indices (an unreplicated table costs 4-5 on these effectively random n).
constexpr int EXP_TAB_N = 64, EXP_TAB_REP = 16, EXP_TAB_BYTES = EXP_TAB_N * EXP_TAB_REP * 8;
__device__ const double EXP_T64[EXP_TAB_N] = {
    1, 1.0108892860517005, 1.0218971486541166, 1.0330248790212284,
    1.0442737824274138, 1.0556451783605572, 1.0671404006768237, 1.0787607977571199,
    1.0905077326652577, 1.1023825833078409, 1.1143867425958924, 1.1265216186082418,
    1.1387886347566916, 1.1511892299529827, 1.1637248587775775, 1.1763969916502812,
    1.189207115002721, 1.2021567314527031, 1.215247359980469, 1.22848053610687,
    1.241857812073484, 1.2553807570246911, 1.2690509571917332, 1.2828700160787783,
    1.2968395546510096, 1.3109612115247644, 1.3252366431597413, 1.3396675240533029,
    1.3542555469368927, 1.3690024229745905, 1.383909881963832, 1.3989796725383112,
    1.4142135623730951, 1.42961333839197, 1.4451808069770467, 1.460917794180647,
    1.4768261459394993, 1.4929077282912648, 1.5091644275934228, 1.5255981507445384,
    1.5422108254079407, 1.5590044002378369, 1.5759808451078865, 1.593142151342267,
    1.6104903319492543, 1.6280274218573478, 1.6457554781539649, 1.6636765803267364,
    1.681792830507429, 1.7001063537185235, 1.7186192981224779, 1.7373338352737062,
    1.7562521603732995, 1.7753764925265212, 1.7947090750031072, 1.8142521755003989,
    1.8340080864093424, 1.8539791250833855, 1.8741676341103, 1.8945759815869656,
    1.9152065613971474, 1.9360617934922943, 1.9571441241754002, 1.9784560263879509,
};
__device__ __forceinline__ double lds_const_f64(uint32_t addr) {   // read-only shared data: free to be scheduled
    double v;
    asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
// tab = shared-space byte address of this lane's copy of the table (table base + 8 (lane & 15))
__device__ __forceinline__ double exp_profile(double x, uint32_t tab) {
    const double xc = (x < -745.0) ? -745.0 : x;
    const double sft = fma(xc, 92.33248261689366, 6755399441055744.0);
    const int n = __double2loint(sft);
    const double nd = sft - 6755399441055744.0;
    double r = fma(nd, -0.010830424696905538, xc);
    r = fma(nd, 6.563929801064195e-13, r);
    const double t = lds_const_f64(tab + (uint32_t)((n & (EXP_TAB_N - 1)) * (EXP_TAB_REP * 8)));
    double q = fma(r, 0.008333339152857227, 0.041666707403338075);
    q = fma(r, q, 0.16666666666664534);
    q = fma(r, q, 0.4999999999998506);
    const double p = fma(r * r, q, r);
    const double scale = __hiloint2double((max(n >> 6, -1023) + 1023) << 20, 0);
    return fma(t, p, t) * scale;
}

__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}

constexpr int SC_DOUBLES = 16;   // one broadcast block of the scalar section
constexpr int LBFGS_M = 10;
enum { CMD_EVAL = 1, CMD_EXIT = 2 };

// one trial point of the line search: step, value, slope, serial number
struct Pt {
    double a, v, s;
    int id;
};
enum { S_P0 = 0, S_LAST, S_PREVOK, S_IA, S_IB, S_A, S_B, S_C, S_NA, S_NB, S_U, S_V, S_W, S_JA, S_JB, S_T1, S_T2, S_NSLOT };
struct LsShared {
    Pt pt[S_NSLOT];
    double x[3], s[3], phi_0, dphi_0, phi_lim;
    int npts, err, ntrials;
};

// Per-CTA state in shared memory (offset 0 of the dynamic shared memory).  Everything the scalar optimiser code keeps
// between evaluations lives here or in registers: the 126 KB of L1 left beside three CTAs' crops cannot hold
// per-thread stacks, and local-memory traffic was what bounded the first version of this kernel (profiles/).
struct EvalShared {
    double sc[SC_DOUBLES];             // broadcast of the dual-number scalar section (master -> everyone)
    double xl[3], fl, gl[3];           // NLSolversBase caches the last evaluated point by value
    double hist_dx[LBFGS_M][3], hist_dg[LBFGS_M][3], hist_rho[LBFGS_M], hist_alpha[LBFGS_M];
    double vx[3], vxp[3], vg[3], vgp[3], vs[3], vdx[3], vdg[3];   // LBFGS iterate, gradient, direction, differences
    LsShared ls;                       // line-search points and base point / direction
    uint32_t crop_s, fbuf_s;           // shared-space byte addresses: staged crop, profile plane (same indexing)
    uint32_t tab_s;                    // ... and the exp table
    int n_a, pitch, TV, m0, rodd, plane, have_last, cmd;
    int newstar, parity;               // first evaluation of a star: the workers look at the crop's mbarrier phase too
    int profile;                       // SHOPT_PROFILE_GAUSSIAN / _KOLMOGOROV
    int sym;                           // pass A over pixel pairs (SHOPT_FIT_SYM=0: every pixel on its own)
};
struct FG {
    double f, g0, g1, g2;
};
extern __shared__ __align__(16) unsigned char smem_raw[];
constexpr int ES_BYTES = (sizeof(EvalShared) + 15) & ~15;
// A CTA serves NSLOT stars at a time ("slots"): each slot has its own master warp, optimiser state, reduction scratch
// and staged crop; the worker warps alternate between the slots, so that one star's serial phase (the master's scalar
// chain between two evaluations: ~40 % of an evaluation's period when a CTA held one star) hides behind the other
// star's pixel passes.  The profile plane is shared: it is live only between the two passes of one evaluation.
template <int THREADS> __host__ __device__ constexpr int slot_stride() { return ES_BYTES + 2 * (THREADS / 32) * 6 * (int)sizeof(double); }
template <int THREADS> __device__ __forceinline__ EvalShared &eval_shared(int slot) {
    return *reinterpret_cast<EvalShared *>(smem_raw + slot * slot_stride<THREADS>());
}
template <int THREADS> __device__ __forceinline__ double *eval_scratch(int slot) {
    return reinterpret_cast<double *>(smem_raw + slot * slot_stride<THREADS>() + ES_BYTES);
}
// the slot a master warp owns (master warps follow the THREADS worker threads)
template <int THREADS> __device__ __forceinline__ int master_slot() { return ((int)threadIdx.x - THREADS) >> 5; }

// Barrier of one evaluation (id 1; the star loop uses __syncthreads = id 0).  The master warp and the worker warps
// arrive from different code, which named barriers allow as long as every warp of the CTA arrives.
// THREADS counts the WORKER threads (warps 0 .. THREADS/32 - 1); the master warp is the extra, last warp of the CTA.
template <int THREADS>
__device__ __forceinline__ void bar_eval(int slot) { asm volatile("bar.sync %0, %1;" ::"r"(1 + slot), "n"(THREADS + 32) : "memory"); }

// Deterministic warp sums with a fixed tree that halves the number of live values at each step instead of carrying
// every value through five butterflies: 6 (8) 64-bit shuffles for 4 (5) values instead of 20 (25).
// Result: value i on lane 8 i (i < 4); the fifth value on lane 4.
__device__ __forceinline__ double warp_reduce4(double v0, double v1, double v2, double v3) {
    const int lane = threadIdx.x & 31;
    const bool b4 = lane & 16, b3 = lane & 8;
    double k0 = (b4 ? v2 : v0) + __shfl_xor_sync(0xffffffffu, b4 ? v0 : v2, 16);
    double k1 = (b4 ? v3 : v1) + __shfl_xor_sync(0xffffffffu, b4 ? v1 : v3, 16);
    double k = (b3 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, b3 ? k0 : k1, 8);
    k += __shfl_xor_sync(0xffffffffu, k, 4);
    k += __shfl_xor_sync(0xffffffffu, k, 2);
    k += __shfl_xor_sync(0xffffffffu, k, 1);
    return k;
}
__device__ __forceinline__ double warp_reduce5(double v0, double v1, double v2, double v3, double v4) {
    const int lane = threadIdx.x & 31;
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    double k0 = (b4 ? v2 : v0) + __shfl_xor_sync(0xffffffffu, b4 ? v0 : v2, 16);
    double k1 = (b4 ? v3 : v1) + __shfl_xor_sync(0xffffffffu, b4 ? v1 : v3, 16);
    double k4 = v4 + __shfl_xor_sync(0xffffffffu, v4, 16);
    double k = (b3 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, b3 ? k0 : k1, 8);
    k4 += __shfl_xor_sync(0xffffffffu, k4, 8);
    k = (b2 ? k4 : k) + __shfl_xor_sync(0xffffffffu, b2 ? k : k4, 4);
    k += __shfl_xor_sync(0xffffffffu, k, 2);
    k += __shfl_xor_sync(0xffffffffu, k, 1);
    return k;
}

// ---------------------------------------------------------------------------------------------
// Kolmogorov profile (radialProfiles.jl:11-24).  Fkolmogorov(k, 1) as written is exp(-C k) with
// C = (24 gamma(6.5) 0.2)^(5/6) / (2 pi)^(5/3), so fkolmogorov's Hankel transform has the closed form
//     f(x) = (1 / 2 pi) C / (C^2 + x)^(3/2),   x = squareRadius = kappa q        (no quadrature)
// and -df/dx = h(x) = 1.5 f / (C^2 + x).  With f replaced by h in the DERIVATIVE sums (d f / d eta = -h d x / d eta where
// the Gaussian has -f d x / d eta) the master's chain rule is unchanged: pass A delivers sum f and the du^2, du dv, dv^2
// moments of h, pass B delivers the cost, sum rho f and the three moments of rho h, in the slots the Gaussian passes use.
// The reference never calls this profile (dead code), so it is a plain two-pass loop: no plane, no unrolling.
// ---------------------------------------------------------------------------------------------
constexpr double KOLMO_C2 = 374.54292214306446;    // C^2, C = 19.353111433127864
constexpr double KOLMO_CN = 3.0801433487905743;    // C / (2 pi)
template <int THREADS>
__device__ __noinline__ void eval_pixels_kolmogorov(int slot) {
    constexpr int NWARPS = THREADS / 32;
    EvalShared &es = eval_shared<THREADS>(slot);
    double *scrA = eval_scratch<THREADS>(slot);
    double *scrB = scrA + NWARPS * 4;
    const double *s = es.sc;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_a = es.n_a, pitch = es.pitch, TV = es.TV, m0 = es.m0, rodd = es.rodd;
    const uint32_t crop_s = es.crop_s;
    const int tv = threadIdx.x / n_a, tu = threadIdx.x - tv * n_a;
    const bool active = tv < TV;
    const double du = (double)(tu + 1) - n_a / 2.0;
    const double dv0 = (double)(tv + 1) - n_a / 2.0;
    const double A11 = s[0], A22 = s[1], G2 = s[2], MS = s[3];
    const double qa = A11 * A11 + G2 * G2, qb = 2.0 * (G2 * (A11 + A22)), qc = G2 * G2 + A22 * A22;
    const double c0 = (MS * qa) * (du * du), c1 = (MS * qb) * du, cc = MS * qc;
    const double step = (double)TV;
    double S0 = 0.0, H0 = 0.0, H1 = 0.0, H2 = 0.0;
    if (active) {
        double dv = dv0;
        for (int v = tv; v < n_a; v += TV) {
            const double t = KOLMO_C2 + fma(dv, fma(cc, dv, c1), c0);
            const double f = KOLMO_CN / (t * sqrt(t)), h = 1.5 * (f / t), hd = h * dv;
            S0 += f;
            H0 += h;
            H1 += hd;
            H2 = fma(hd, dv, H2);
            dv += step;
        }
    }
    {
        const double k = warp_reduce4(S0, (du * du) * H0, du * H1, H2);
        if ((lane & 7) == 0) scrA[warp * 4 + (lane >> 3)] = k;
    }
    bar_eval<THREADS>(slot);
    double sumf = scrA[0];
#pragma unroll
    for (int w = 1; w < NWARPS; w++) sumf += scrA[w * 4];
    const double A = 1.0 / sumf;
    double R0 = 0.0, P0 = 0.0, Q0 = 0.0, Q1 = 0.0, Q2 = 0.0;
    if (active) {
        double dv = dv0;
        for (int v = tv; v < n_a; v += TV) {
            const uint32_t o = (uint32_t)((v * pitch + ((m0 + v * rodd) & 1) + tu) * 8);
            const double d = lds_f64(crop_s + o);
            if (!isnan(d)) {
                const double t = KOLMO_C2 + fma(dv, fma(cc, dv, c1), c0);
                const double f = KOLMO_CN / (t * sqrt(t)), h = 1.5 * (f / t);
                const double rho = fma(A, f, -d);
                const double th = rho * h, td = th * dv;
                R0 = fma(rho, rho, R0);
                P0 = fma(rho, f, P0);
                Q0 += th;
                Q1 += td;
                Q2 = fma(td, dv, Q2);
            }
            dv += step;
        }
    }
    {
        const double k = warp_reduce5(P0, (du * du) * Q0, du * Q1, Q2, 0.5 * R0);
        if ((lane & 7) == 0) scrB[warp * 5 + 1 + (lane >> 3)] = k;
        else if (lane == 4) scrB[warp * 5] = k;
    }
    bar_eval<THREADS>(slot);
}

// The pixel passes of one evaluation, executed by the worker warps between the evaluation's first barrier (value block
// published by the master) and its third.  The totals are needed by the master alone; a worker needs sum f (the
// normalisation) for the second pass.
template <int THREADS>
__device__ __forceinline__ void eval_pixels(int slot) {
    constexpr int NWARPS = THREADS / 32;
    constexpr int UA = 8;                                             // pixels in flight per thread in pass A
    EvalShared &es = eval_shared<THREADS>(slot);
    double *scrA = eval_scratch<THREADS>(slot);                       // NWARPS x 4
    double *scrB = scrA + NWARPS * 4;                                  // NWARPS x 5
    const double *s = es.sc;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_a = es.n_a, pitch = es.pitch, TV = es.TV, m0 = es.m0, rodd = es.rodd;
    const bool plane = es.plane != 0;
    const uint32_t crop_s = es.crop_s, fbuf_s = es.fbuf_s, tab = es.tab_s + 8u * (lane & 15);
    const int tv = threadIdx.x / n_a, tu = threadIdx.x - tv * n_a;
    const bool active = tv < TV;
    const double du = (double)(tu + 1) - n_a / 2.0;    // analyticLBFGS.jl:28,41: the centre is (r/2, c/2) of the crop
    const double dv0 = (double)(tv + 1) - n_a / 2.0;
    const double A11 = s[0], A22 = s[1], G2 = s[2], MS = s[3];
    // q = qa du^2 + qb du dv + qc dv^2
    const double qa = A11 * A11 + G2 * G2, qb = 2.0 * (G2 * (A11 + A22)), qc = G2 * G2 + A22 * A22;
    const double c0 = (MS * qa) * (du * du), c1 = (MS * qb) * du, cc = MS * qc;
    const double step = (double)TV;
    const bool tame = MS >= 0.0 && MS <= 1e300;        // else (never for a finite profile): the library's exp, simple loops
    // byte offset of pixel (tu, v) in the staged crop / the profile plane: (v pitch + ((m0 + v rodd) & 1) + tu) 8; along
    // v -> v + TV it advances by TV pitch 8, plus or minus 8 when the column's alignment bit flips (odd r and odd TV)
    const uint32_t o0 = (uint32_t)((tv * pitch + ((m0 + tv * rodd) & 1) + tu) * 8);
    const int flip = (TV & rodd & 1) ? 8 : 0;
    const int inc0 = TV * pitch * 8 + ((((m0 + tv * rodd) & 1) == 0) ? flip : -flip), incsum = 2 * TV * pitch * 8;
    // pass A: profile moments over ALL pixels (analyticLBFGS.jl:24-34)
    double S0 = 0.0, S1 = 0.0, S2 = 0.0;
    double E0 = 0.0, E1 = 0.0, E2 = 0.0, E3 = 0.0;                 // pixels this thread serves outside its own row (below)
    if (tame && plane && es.sym) {
        // The profile is symmetric under the point reflection about its centre (n_a/2, n_a/2):
        //     pixel (ui, vi) <-> (n_a-2-ui, n_a-2-vi)  has  du -> -du, dv -> -dv,  c1 -> -c1,  hence the SAME exponent bits,
        // and the products the moments are made of (f, f du^2, f du dv, f dv^2) are even under it.  So the exp of a
        // pixel pair is evaluated once, stored to both places of the plane (pass B needs f everywhere: the data are not
        // symmetric) and counted twice.  Only row n_a-1 and column n_a-1 have no partner inside the window: those
        // 2 n_a - 1 pixels are shared out one per thread, whatever row the thread owns, and enter the reduction with
        // their own du, dv.  Pass A drops from n_a^2 to ~n_a^2 / 2 + 2 n_a exps and moment updates (41.9 -> 37.8 ms per
        // 100k stars).  The same pairing of pass B ((rho + rho') f per pair, one load of f) was measured SLOWER: 39.2 ms.
        const int nm2 = n_a - 2;
        if (active && tu < n_a - 1) {
            const int up = nm2 - tu;                                  // partner row
            double T0 = 0.0, T1 = 0.0, T2 = 0.0;
            int vi = tv;
            double dv = dv0;
            for (; 2 * (vi + 3 * TV) < nm2; vi += 4 * TV) {           // four pairs at a time
                double dvj[4], fv[4];
#pragma unroll
                for (int j = 0; j < 4; j++) { dvj[j] = dv; dv += step; }
#pragma unroll
                for (int j = 0; j < 4; j++) fv[j] = exp_profile(-fma(dvj[j], fma(cc, dvj[j], c1), c0), tab);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int v1 = vi + j * TV, v2 = nm2 - v1;
                    sts_f64(fbuf_s + (uint32_t)((v1 * pitch + ((m0 + v1 * rodd) & 1) + tu) * 8), fv[j]);
                    sts_f64(fbuf_s + (uint32_t)((v2 * pitch + ((m0 + v2 * rodd) & 1) + up) * 8), fv[j]);
                    const double fd = fv[j] * dvj[j];
                    T0 += fv[j];
                    T1 += fd;
                    T2 = fma(fd, dvj[j], T2);
                }
            }
            for (; 2 * vi < nm2; vi += TV, dv += step) {
                const double fv = exp_profile(-fma(dv, fma(cc, dv, c1), c0), tab);
                const int v2 = nm2 - vi;
                sts_f64(fbuf_s + (uint32_t)((vi * pitch + ((m0 + vi * rodd) & 1) + tu) * 8), fv);
                sts_f64(fbuf_s + (uint32_t)((v2 * pitch + ((m0 + v2 * rodd) & 1) + up) * 8), fv);
                const double fd = fv * dv;
                T0 += fv;
                T1 += fd;
                T2 = fma(fd, dv, T2);
            }
            S0 = 2.0 * T0; S1 = 2.0 * T1; S2 = 2.0 * T2;
            if (2 * vi == nm2 && tu <= up) {                          // the column that is its own partner (n_a even): rows pair up
                const double fv = exp_profile(-fma(dv, fma(cc, dv, c1), c0), tab);
                const double w = (tu == up) ? 1.0 : 2.0;
                sts_f64(fbuf_s + (uint32_t)((vi * pitch + ((m0 + vi * rodd) & 1) + tu) * 8), fv);
                if (tu != up) sts_f64(fbuf_s + (uint32_t)((vi * pitch + ((m0 + vi * rodd) & 1) + up) * 8), fv);
                const double fd = (w * fv) * dv;
                S0 += w * fv;
                S1 += fd;
                S2 = fma(fd, dv, S2);
            }
        }
        const int e = threadIdx.x;
        if (e < 2 * n_a - 1) {                                        // the unpaired row and column
            const int ui = e < n_a ? n_a - 1 : e - n_a, vi = e < n_a ? e : n_a - 1;
            const double pdu = (double)(ui + 1) - n_a / 2.0, pdv = (double)(vi + 1) - n_a / 2.0;
            const double pc0 = (MS * qa) * (pdu * pdu), pc1 = (MS * qb) * pdu;
            const double fv = exp_profile(-fma(pdv, fma(cc, pdv, pc1), pc0), tab);
            sts_f64(fbuf_s + (uint32_t)((vi * pitch + ((m0 + vi * rodd) & 1) + ui) * 8), fv);
            E0 = fv; E1 = fv * (pdu * pdu); E2 = fv * (pdu * pdv); E3 = fv * (pdv * pdv);
        }
    } else if (active) {
        double dv = dv0;
        uint32_t o = o0;
        int inc = inc0;
        int v = tv;
        for (; tame && v + (UA - 1) * TV < n_a; v += UA * TV) {    // UA pixels at a time: independent exp chains interleave
            double dvj[UA], fv[UA];
            uint32_t oj[UA];
#pragma unroll
            for (int j = 0; j < UA; j++) {
                dvj[j] = dv; oj[j] = o;
                dv += step; o += inc; inc = incsum - inc;
            }
#pragma unroll
            for (int j = 0; j < UA; j++) fv[j] = exp_profile(-fma(dvj[j], fma(cc, dvj[j], c1), c0), tab);
#pragma unroll
            for (int j = 0; j < UA; j++) {
                if (plane) sts_f64(fbuf_s + oj[j], fv[j]);
                const double fd = fv[j] * dvj[j];
                S0 += fv[j];
                S1 += fd;
                S2 = fma(fd, dvj[j], S2);
            }
        }
        for (; v < n_a; v += TV) {
            const double xa = -fma(dv, fma(cc, dv, c1), c0);
            const double fv = tame ? exp_profile(xa, tab) : exp(xa);
            if (plane) sts_f64(fbuf_s + o, fv);
            o += inc; inc = incsum - inc;
            const double fd = fv * dv;
            S0 += fv;
            S1 += fd;
            S2 = fma(fd, dv, S2);
            dv += step;
        }
    }
    {   // sum f, sum f du^2, sum f du dv, sum f dv^2
        const double k = warp_reduce4(S0 + E0, (du * du) * S0 + E1, du * S1 + E2, S2 + E3);
        if ((lane & 7) == 0) scrA[warp * 4 + (lane >> 3)] = k;
    }
    bar_eval<THREADS>(slot);
    double sumf = scrA[0];
#pragma unroll
    for (int w = 1; w < NWARPS; w++) sumf += scrA[w * 4];
    const double A = 1.0 / sumf;
    // pass B: residual moments over non-NaN pixels (analyticLBFGS.jl:36-44)
    double R0 = 0.0, R1 = 0.0, R2 = 0.0, R3 = 0.0;
    if (active) {
        double dv = dv0;
        uint32_t o = o0;
        int inc = inc0;
        int v = tv;
        if (plane) {
            for (; v + 3 * TV < n_a; v += 4 * TV) {    // four pixels at a time: the loads are issued together
                double dvj[4], d[4], fv[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    dvj[j] = dv;
                    d[j] = lds_f64(crop_s + o); fv[j] = lds_f64(fbuf_s + o);
                    dv += step; o += inc; inc = incsum - inc;
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (!isnan(d[j])) {
                        const double rho = fma(A, fv[j], -d[j]);
                        const double t = rho * fv[j], td = t * dvj[j];
                        R0 = fma(rho, rho, R0);
                        R1 += t;
                        R2 += td;
                        R3 = fma(td, dvj[j], R3);
                    }
                }
            }
        }
        for (; v < n_a; v += TV) {
            const double d = lds_f64(crop_s + o);
            if (!isnan(d)) {
                const double xa = -fma(dv, fma(cc, dv, c1), c0);
                const double fv = plane ? lds_f64(fbuf_s + o) : (tame ? exp_profile(xa, tab) : exp(xa));
                const double rho = fma(A, fv, -d);
                const double t = rho * fv, td = t * dv;
                R0 = fma(rho, rho, R0);
                R1 += t;
                R2 += td;
                R3 = fma(td, dv, R3);
            }
            o += inc; inc = incsum - inc;
            dv += step;
        }
    }
    {   // cost, sum rho f, its du^2, du dv, dv^2 moments
        const double k = warp_reduce5(R1, (du * du) * R1, du * R2, R3, 0.5 * R0);
        if ((lane & 7) == 0) scrB[warp * 5 + 1 + (lane >> 3)] = k;
        else if (lane == 4) scrB[warp * 5] = k;
    }
    bar_eval<THREADS>(slot);
}

// value_gradient!(d, x): the reference runs g! (ForwardDiff over cost) and then cost(x); the value channel of the dual
// pass is the same operation sequence, so one evaluation yields both.  MASTER WARP ONLY: it owns the optimiser state
// and does no pixel work.  It publishes the VALUES of the scalar section (all the pixel passes need), releases the
// workers, and evaluates the dual-number partials (a dozen dependent IEEE divisions and square roots) while they run.
template <int THREADS>
__device__ __noinline__ FG eval_fg(double x0, double x1, double x2) {
    constexpr int NWARPS = THREADS / 32;
    const int slot = master_slot<THREADS>();
    EvalShared &es = eval_shared<THREADS>(slot);
    FG out;
    if (es.have_last && x0 == es.xl[0] && x1 == es.xl[1] && x2 == es.xl[2]) {
        out.f = es.fl; out.g0 = es.gl[0]; out.g1 = es.gl[1]; out.g2 = es.gl[2];
        return out;
    }
    double *s = es.sc;
    const int lane = threadIdx.x & 31;
    // value channel of the scalar section: all the pixel passes need
    const ScalarVals sv = scalar_values(x0, x1, x2, lane);
    const double A11 = add_(1.0, sv.g1), A22 = sub_(1.0, sv.g1), G2 = sv.g2, MS = sv.ms;
    if (lane == 0) {
        s[0] = A11; s[1] = A22; s[2] = G2; s[3] = MS;
        es.cmd = CMD_EVAL;
    }
    bar_eval<THREADS>(slot);                               // workers start the pixel passes
    // partial channel of this lane (sigma, e1, e2 on lanes 0, 1, 2), overlapped with pass A
    double a11_d, g2_d, ms_d;
    scalar_partials(sv, x0, x1, x2, lane, a11_d, g2_d, ms_d);
    const double a22_d = -a11_d;
    const double *scrA = eval_scratch<THREADS>(slot);
    const double *scrB = scrA + NWARPS * 4;
    double M[4], T[5];
    bar_eval<THREADS>(slot);                               // pass A partials are in
#pragma unroll
    for (int i = 0; i < 4; i++) {
        double t = scrA[i];
#pragma unroll
        for (int w = 1; w < NWARPS; w++) t += scrA[w * 4 + i];
        M[i] = t;
    }
    const double qa = A11 * A11 + G2 * G2, qb = 2.0 * (G2 * (A11 + A22)), qc = G2 * G2 + A22 * A22;
    const double A = 1.0 / M[0];
    // rebuild the reference's sums from the moments, then the chain rule, one channel per lane:
    //   sum.d_j = -(ms.d_j Sq + 2 ms (a11.d_j S2 + g2.d_j S3 + a22.d_j S4))
    const double Sq = qa * M[1] + qb * M[2] + qc * M[3];
    const double Sw1u = A11 * M[1] + G2 * M[2];
    const double Sx = (A11 + A22) * M[2] + G2 * (M[1] + M[3]);
    const double Sw2v = G2 * M[2] + A22 * M[3];
    const double dS = -(ms_d * Sq + 2.0 * MS * (a11_d * Sw1u + g2_d * Sx + a22_d * Sw2v));
    const double dA = -(A / M[0]) * dS;
    bar_eval<THREADS>(slot);                               // pass B partials are in
#pragma unroll
    for (int i = 0; i < 5; i++) {
        double t = scrB[i];
#pragma unroll
        for (int w = 1; w < NWARPS; w++) t += scrB[w * 5 + i];
        T[i] = t;
    }
    const double Tq = qa * T[2] + qb * T[3] + qc * T[4];
    const double Tw1u = A11 * T[2] + G2 * T[3];
    const double Tx = (A11 + A22) * T[3] + G2 * (T[2] + T[4]);
    const double Tw2v = G2 * T[3] + A22 * T[4];
    const double dT = -(ms_d * Tq + 2.0 * MS * (a11_d * Tw1u + g2_d * Tx + a22_d * Tw2v));
    const double gl = dA * T[1] + A * dT;                  // this lane's gradient component
    double g[3];
    g[0] = bcast_(gl, 0); g[1] = bcast_(gl, 1); g[2] = bcast_(gl, 2);
    out.f = T[0]; out.g0 = g[0]; out.g1 = g[1]; out.g2 = g[2];
    // every lane of the master warp stores the same bits
    es.xl[0] = x0; es.xl[1] = x1; es.xl[2] = x2;
    es.fl = out.f; es.gl[0] = g[0]; es.gl[1] = g[1]; es.gl[2] = g[2];
    es.have_last = 1;
    es.newstar = 0;
    __syncwarp();
    return out;
}

// Worker warps: serve the slots in strict alternation -- wait for slot k's master (its next evaluation request), run
// the two pixel passes, move on to the other slot -- until every master has said it is out of stars.  Between two stars
// a master simply stays away from its barrier; the workers meanwhile serve the other slot and then wait.
template <int THREADS, int NSLOT>
__device__ __forceinline__ void worker_loop(uint64_t *bars) {
    int alive = (1 << NSLOT) - 1;
    for (int k = 0; alive; k = (k + 1 == NSLOT) ? 0 : k + 1) {
        if (!(alive & (1 << k))) continue;
        bar_eval<THREADS>(k);
        EvalShared &es = eval_shared<THREADS>(k);
        const int cmd = *reinterpret_cast<volatile int *>(&es.cmd);
        if (cmd == CMD_EXIT) { alive &= ~(1 << k); continue; }
        if (es.newstar) mbar_wait(&bars[k], (uint32_t)es.parity);   // the crop's bulk copies, observed by this thread too
        if (es.profile) eval_pixels_kolmogorov<THREADS>(k);
        else eval_pixels<THREADS>(k);
    }
}

// ---------------------------------------------------------------------------------------------
// HagerZhang line search (LineSearches 7.2.0 defaults), value form: trial points are
// carried as (alpha, phi, dphi, id) tuples instead of indices into a history array.  The only use the
// algorithm makes of older history is "the most recent earlier point with phi <= phi_lim" during
// bracketing, which is tracked incrementally.
// ---------------------------------------------------------------------------------------------
enum { HZ_OK = 0, HZ_LSEXC = 1, HZ_ERROR = 2 };

__device__ __forceinline__ double eps_of(double b) {   // Julia eps(x::Float64)
    double a = fabs(b);
    if (!isfinite(a)) return nan("");
    if (a < DBL_MIN) return 4.9406564584124654e-324;
    int e;
    (void)frexp(a, &e);
    return ldexp(DBL_EPSILON, e - 1);
}
__device__ __forceinline__ double nextfloat_(double x) { return nextafter(x, (double)INFINITY); }

// The trial points live in SHARED memory slots and the routines pass slot numbers, like the index-into-history style
// of LineSearches.jl itself.  eval_fg is a real call (one copy of the evaluator instead of one per call site), and every
// value that is live in registers across a call is saved to the local stack by the ABI: with the points in registers
// that was 20 % of the master's instructions, each an L1 miss (the L1 left beside three CTAs' crops is a few KB).

template <int THREADS>
struct LineSearch {
    __device__ __forceinline__ static void phidphi(LsShared &L, double alpha, double &phi, double &dphi) {
        const FG r = eval_fg<THREADS>(add_(L.x[0], mul_(alpha, L.s[0])), add_(L.x[1], mul_(alpha, L.s[1])),
                                      add_(L.x[2], mul_(alpha, L.s[2])));
        phi = r.f;
        dphi = add_(add_(mul_(r.g0, L.s[0]), mul_(r.g1, L.s[1])), mul_(r.g2, L.s[2]));
        L.ntrials++;
    }
    __device__ __forceinline__ static void push(LsShared &L, int slot, double a, double v, double sl) {
        L.pt[slot].a = a; L.pt[slot].v = v; L.pt[slot].s = sl; L.pt[slot].id = L.npts;
        L.npts++;
    }
    __device__ __forceinline__ static void copy(LsShared &L, int dst, int src) {
        if (dst != src) L.pt[dst] = L.pt[src];
    }
    __device__ __forceinline__ static bool wolfe(const LsShared &L, double c, double phi_c, double dphi_c) {
        const double delta = 0.1, sigma = 0.9;
        bool w1 = (mul_(delta, L.dphi_0) >= div_(sub_(phi_c, L.phi_0), c)) && (dphi_c >= mul_(sigma, L.dphi_0));
        bool w2 = (mul_(sub_(mul_(2.0, delta), 1.0), L.dphi_0) >= dphi_c) && (dphi_c >= mul_(sigma, L.dphi_0)) &&
                  (phi_c <= L.phi_lim);
        return w1 || w2;
    }
    // HZ stage U3 (theta = 1/2); inputs ia, ib, outputs oa, ob (slots; outputs may alias inputs)
    __device__ static void bisect(LsShared &L, int ia, int ib, int oa, int ob) {
        copy(L, S_T1, ia); copy(L, S_T2, ib);
        {
            const Pt &A = L.pt[S_T1], &B = L.pt[S_T2];
            if (!(A.s < 0.0) || !(A.v <= L.phi_lim) || !(B.s < 0.0) || !(B.v > L.phi_lim) || !(B.a > A.a)) {
                L.err = 1; copy(L, oa, S_T1); copy(L, ob, S_T2); return;
            }
        }
        while (sub_(L.pt[S_T2].a, L.pt[S_T1].a) > eps_of(L.pt[S_T2].a)) {
            const double d = div_(add_(L.pt[S_T1].a, L.pt[S_T2].a), 2.0);
            double phi_d, g;
            phidphi(L, d, phi_d, g);
            if (!(isfinite(phi_d) && isfinite(g))) { L.err = 1; break; }
            if (g >= 0.0) { copy(L, oa, S_T1); push(L, ob, d, phi_d, g); return; }
            if (phi_d <= L.phi_lim) push(L, S_T1, d, phi_d, g); else push(L, S_T2, d, phi_d, g);
        }
        copy(L, oa, S_T1); copy(L, ob, S_T2);
    }
    // HZ stages U0-U3; outputs must not alias the inputs
    __device__ static void update(LsShared &L, int ia, int ib, int ic, int oa, int ob) {
        const double a = L.pt[ia].a, b = L.pt[ib].a;
        if (!(L.pt[ia].s < 0.0) || !(L.pt[ia].v <= L.phi_lim) || !(L.pt[ib].s >= 0.0) || !(b > a)) {
            L.err = 1; copy(L, oa, ia); copy(L, ob, ib); return;
        }
        const double ca = L.pt[ic].a;
        if (ca < a || ca > b) { copy(L, oa, ia); copy(L, ob, ib); return; }
        if (L.pt[ic].s >= 0.0) { copy(L, oa, ia); copy(L, ob, ic); return; }
        if (L.pt[ic].v <= L.phi_lim) { copy(L, oa, ic); copy(L, ob, ib); return; }
        bisect(L, ia, ic, oa, ob);
    }
    __device__ __forceinline__ static double secant(double a, double b, double dphi_a, double dphi_b) {
        return div_(sub_(mul_(a, dphi_b), mul_(b, dphi_a)), sub_(dphi_b, dphi_a));
    }
    // HZ stages S1-S4 on (S_IA, S_IB) -> (S_A, S_B); returns iswolfe
    __device__ static bool secant2(LsShared &L) {
        copy(L, S_A, S_IA); copy(L, S_B, S_IB);
        if (!(L.pt[S_IA].s < 0.0 && L.pt[S_IB].s >= 0.0)) { L.err = 1; return false; }
        double c = secant(L.pt[S_IA].a, L.pt[S_IB].a, L.pt[S_IA].s, L.pt[S_IB].s);
        if (!isfinite(c)) { L.err = 1; return false; }
        double phi_c, dphi_c;
        phidphi(L, c, phi_c, dphi_c);
        if (!(isfinite(phi_c) && isfinite(dphi_c))) { L.err = 1; return false; }
        push(L, S_W, c, phi_c, dphi_c);
        if (wolfe(L, c, phi_c, dphi_c)) { copy(L, S_A, S_W); copy(L, S_B, S_W); return true; }
        update(L, S_IA, S_IB, S_W, S_U, S_V);
        if (L.err) return false;
        const double a = L.pt[S_U].a, b = L.pt[S_V].a;
        const bool movedB = (L.pt[S_V].id == L.pt[S_W].id), movedA = (L.pt[S_U].id == L.pt[S_W].id);
        if (movedB) c = secant(L.pt[S_IB].a, L.pt[S_V].a, L.pt[S_IB].s, L.pt[S_V].s);
        else if (movedA) c = secant(L.pt[S_IA].a, L.pt[S_U].a, L.pt[S_IA].s, L.pt[S_U].s);
        if ((movedA || movedB) && a <= c && c <= b) {
            phidphi(L, c, phi_c, dphi_c);
            if (!(isfinite(phi_c) && isfinite(dphi_c))) { L.err = 1; return false; }
            push(L, S_W, c, phi_c, dphi_c);
            if (wolfe(L, c, phi_c, dphi_c)) { copy(L, S_A, S_W); copy(L, S_B, S_W); return true; }
            update(L, S_U, S_V, S_W, S_JA, S_JB);
            if (L.err) return false;
            copy(L, S_U, S_JA); copy(L, S_V, S_JB);
        }
        copy(L, S_A, S_U); copy(L, S_B, S_V);
        return false;
    }

    // returns HZ_*; alpha_out = accepted step.  L.x, L.s hold the base point and the direction.
    __device__ static int run(LsShared &L, double c, double phi0, double dphi0, double &alpha_out) {
        const double rho = 5.0, epsilon = 1e-6, gamma = 0.66, psi3 = 0.1;
        const int linesearchmax = 50, iterfinitemax = 52;
        double alphamax = (double)INFINITY;
        alpha_out = 0.0;
        L.phi_0 = phi0; L.dphi_0 = dphi0;
        if (!(isfinite(phi0) && isfinite(dphi0))) return HZ_LSEXC;
        if (dphi0 >= mul_(DBL_EPSILON, fabs(phi0))) return HZ_LSEXC;
        else if (dphi0 >= 0.0) return HZ_OK;
        L.npts = 0; L.err = 0;
        push(L, S_P0, 0.0, phi0, dphi0);
        L.phi_lim = add_(phi0, mul_(epsilon, fabs(phi0)));
        if (!(c >= 0.0)) return HZ_ERROR;
        if (c <= DBL_EPSILON) return HZ_OK;
        if (!(isfinite(c) && c <= alphamax)) return HZ_ERROR;
        double phi_c, dphi_c;
        phidphi(L, c, phi_c, dphi_c);
        int iterfinite = 1;
        while (!(isfinite(phi_c) && isfinite(dphi_c)) && iterfinite < iterfinitemax) {
            iterfinite++;
            c = mul_(c, psi3);
            phidphi(L, c, phi_c, dphi_c);
        }
        if (!(isfinite(phi_c) && isfinite(dphi_c))) return HZ_OK;
        push(L, S_LAST, c, phi_c, dphi_c);
        // most recent point with phi <= phi_lim among the points pushed BEFORE `last`
        copy(L, S_PREVOK, S_P0);
        // bracketing, HZ stages B0-B3
        bool isbracketed = false;
        copy(L, S_IA, S_P0); copy(L, S_IB, S_LAST);
        int iter = 1;
        double cold = -1.0;
        while (!isbracketed && iter < linesearchmax) {
            if (dphi_c >= 0.0) {
                copy(L, S_IB, S_LAST); copy(L, S_IA, S_PREVOK);
                isbracketed = true;
            } else if (L.pt[S_LAST].v > L.phi_lim) {
                bisect(L, S_P0, S_LAST, S_NA, S_NB);
                if (L.err) return HZ_ERROR;
                copy(L, S_IA, S_NA); copy(L, S_IB, S_NB);
                isbracketed = true;
            } else {
                cold = c;
                if (nextfloat_(cold) >= alphamax) { alpha_out = cold; return HZ_OK; }
                c = mul_(c, rho);
                if (c > alphamax) c = alphamax;
                phidphi(L, c, phi_c, dphi_c);
                iterfinite = 1;
                while (!(isfinite(phi_c) && isfinite(dphi_c)) && c > nextfloat_(cold) && iterfinite < iterfinitemax) {
                    alphamax = c;
                    iterfinite++;
                    c = div_(add_(cold, c), 2.0);
                    phidphi(L, c, phi_c, dphi_c);
                }
                if (!(isfinite(phi_c) && isfinite(dphi_c))) { alpha_out = cold; return HZ_OK; }
                if (L.pt[S_LAST].v <= L.phi_lim) copy(L, S_PREVOK, S_LAST);
                push(L, S_LAST, c, phi_c, dphi_c);
            }
            iter++;
        }
        while (iter < linesearchmax) {
            const double a = L.pt[S_IA].a, b = L.pt[S_IB].a;
            if (!(b > a)) return HZ_ERROR;
            if (sub_(b, a) <= eps_of(b)) { alpha_out = a; return HZ_OK; }
            const bool iswolfe = secant2(L);
            if (L.err) return HZ_ERROR;
            if (iswolfe) { alpha_out = L.pt[S_A].a; return HZ_OK; }
            const double A = L.pt[S_A].a, B = L.pt[S_B].a;
            if (!(B > A)) return HZ_ERROR;
            if (sub_(B, A) < mul_(gamma, sub_(b, a))) {
                if (nextfloat_(L.pt[S_IA].v) >= L.pt[S_IB].v && nextfloat_(L.pt[S_A].v) >= L.pt[S_B].v) { alpha_out = A; return HZ_OK; }
                copy(L, S_IA, S_A); copy(L, S_IB, S_B);
            } else {
                c = div_(add_(A, B), 2.0);
                phidphi(L, c, phi_c, dphi_c);
                if (!(isfinite(phi_c) && isfinite(dphi_c))) return HZ_ERROR;
                push(L, S_C, c, phi_c, dphi_c);
                update(L, S_A, S_B, S_C, S_NA, S_NB);
                if (L.err) return HZ_ERROR;
                copy(L, S_IA, S_NA); copy(L, S_IB, S_NB);
            }
            iter++;
        }
        alpha_out = L.pt[S_IA].a;
        return HZ_LSEXC;
    }
};

// ---------------------------------------------------------------------------------------------
// Optim 1.7.8 LBFGS (m = 10, InitialStatic(alpha = 1), HagerZhang, scaleinvH0, no preconditioner) under
// the generic optimize() loop -- SURVEY.md A.1.  Scalar work, executed redundantly by every thread.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double dot3(const double a[3], const double b[3]) {
    return add_(add_(mul_(a[0], b[0]), mul_(a[1], b[1])), mul_(a[2], b[2]));
}
__device__ __forceinline__ double maxabs3(const double a[3]) { return fmax(fmax(fabs(a[0]), fabs(a[1])), fabs(a[2])); }

template <int THREADS>
__device__ int lbfgs_fit(const double x0[3], double g_tol, int max_iter, double out_p[3], int &iters_out) {
    constexpr int M = LBFGS_M;
    EvalShared &es = eval_shared<THREADS>(master_slot<THREADS>());
    // the vectors of the iteration live in shared memory (every lane of the master warp stores the same bits): nothing
    // but a few scalars is live in registers across the evaluator calls
    double (&x)[3] = es.vx, (&x_prev)[3] = es.vxp, (&g)[3] = es.vg, (&g_prev)[3] = es.vgp, (&s)[3] = es.vs,
           (&dx)[3] = es.vdx, (&dg)[3] = es.vdg;
    x[0] = x0[0]; x[1] = x0[1]; x[2] = x0[2];
    LsShared &L = es.ls;
    L.ntrials = 0;
    double (*dxh)[3] = es.hist_dx, (*dgh)[3] = es.hist_dg;
    double *rho = es.hist_rho, *tl_alpha = es.hist_alpha;
    double f, f_prev = nan("");
    int pseudo = 0, iteration = 0, counter_f_tol = 0, status;
    { const FG r0 = eval_fg<THREADS>(x[0], x[1], x[2]); f = r0.f; g[0] = r0.g0; g[1] = r0.g1; g[2] = r0.g2; }
    const bool g_conv0 = maxabs3(g) <= g_tol;
    const bool stopped0 = !isfinite(f) || !isfinite(g[0]) || !isfinite(g[1]) || !isfinite(g[2]);
    bool converged = g_conv0 || stopped0;
    status = stopped0 ? SHOPT_FIT_NONFINITE : SHOPT_FIT_CONVERGED;
    bool broke = false;
    while (!converged && iteration < max_iter) {
        iteration++;
        pseudo++;
        {   // two-loop recursion
            const int lower = pseudo - M, upper = pseudo - 1;
            double q[3] = {g[0], g[1], g[2]};
            for (int index = upper; index >= lower; index--) {
                if (index < 1) continue;
                const int i = (index - 1) % M;
                tl_alpha[i] = mul_(rho[i], dot3(dxh[i], q));
#pragma unroll
                for (int k = 0; k < 3; k++) q[k] = sub_(q[k], mul_(tl_alpha[i], dgh[i][k]));
            }
            if (pseudo > 1) {
                const int i = (upper - 1) % M;
                const double scaling = div_(dot3(dxh[i], dgh[i]), dot3(dgh[i], dgh[i]));
#pragma unroll
                for (int k = 0; k < 3; k++) s[k] = mul_(scaling, q[k]);
            } else {
#pragma unroll
                for (int k = 0; k < 3; k++) s[k] = q[k];
            }
            for (int index = lower; index <= upper; index++) {
                if (index < 1) continue;
                const int i = (index - 1) % M;
                const double beta = mul_(rho[i], dot3(dgh[i], s));
#pragma unroll
                for (int k = 0; k < 3; k++) s[k] = add_(s[k], mul_(dxh[i][k], sub_(tl_alpha[i], beta)));
            }
#pragma unroll
            for (int k = 0; k < 3; k++) s[k] = mul_(-1.0, s[k]);
        }
#pragma unroll
        for (int k = 0; k < 3; k++) g_prev[k] = g[k];
        double dphi_0 = dot3(g, s);
        if (dphi_0 >= 0.0) {   // reset_search_direction!
            pseudo = 1;
#pragma unroll
            for (int k = 0; k < 3; k++) s[k] = -g[k];
            dphi_0 = dot3(g, s);
        }
        const double phi_0 = f;
        f_prev = phi_0;
#pragma unroll
        for (int k = 0; k < 3; k++) { x_prev[k] = x[k]; L.x[k] = x[k]; L.s[k] = s[k]; }
        double alpha = 1.0;   // InitialStatic(alpha = 1.0, scaled = false)
        const int hz = LineSearch<THREADS>::run(L, alpha, phi_0, dphi_0, alpha);
        if (hz == HZ_ERROR) { status = SHOPT_FIT_THREW; broke = true; break; }
#pragma unroll
        for (int k = 0; k < 3; k++) { dx[k] = mul_(alpha, s[k]); x[k] = add_(x[k], dx[k]); }
        if (hz == HZ_LSEXC) { status = SHOPT_FIT_LS_FAILED; broke = true; break; }
        { const FG r1 = eval_fg<THREADS>(x[0], x[1], x[2]); f = r1.f; g[0] = r1.g0; g[1] = r1.g1; g[2] = r1.g2; }
        // assess_convergence with x_abstol = x_reltol = f_abstol = f_reltol = 0
        const double xchg = fmax(fmax(fabs(sub_(x[0], x_prev[0])), fabs(sub_(x[1], x_prev[1]))), fabs(sub_(x[2], x_prev[2])));
        const bool x_conv = xchg <= 0.0;
        const bool f_conv = fabs(sub_(f, f_prev)) <= 0.0;
        const bool g_conv = maxabs3(g) <= g_tol;
        counter_f_tol = f_conv ? counter_f_tol + 1 : 0;
        converged = x_conv || g_conv || (counter_f_tol > 1);
        // update_h!
#pragma unroll
        for (int k = 0; k < 3; k++) dg[k] = sub_(g[k], g_prev[k]);
        const double rho_it = div_(1.0, dot3(dx, dg));
        if (isinf(rho_it)) {
            pseudo = 0;
        } else {
            const int idx = (pseudo - 1) % M;
#pragma unroll
            for (int k = 0; k < 3; k++) { dxh[idx][k] = dx[k]; dgh[idx][k] = dg[k]; }
            rho[idx] = rho_it;
        }
        if (!isfinite(g[0]) || !isfinite(g[1]) || !isfinite(g[2])) { status = SHOPT_FIT_NONFINITE; broke = true; break; }
    }
    if (!broke && !(g_conv0 || stopped0)) status = converged ? SHOPT_FIT_CONVERGED : SHOPT_FIT_MAXITER;
    if (status == SHOPT_FIT_THREW) {
        out_p[0] = out_p[1] = out_p[2] = 0.0;   // shopt.jl:155-162 catch
    } else {
        out_p[0] = mul_(x[0], x[0]);
        e_to_g(x[1], x[2], out_p[1], out_p[2]);
    }
    iters_out = iteration;
    return status;
}

// ---------------------------------------------------------------------------------------------
// Kernel: persistent CTAs pull star indices from a global counter; crop staging by bulk copies.
// Shared memory layout (dynamic): [ mbarriers 2 x 8 B | pad to 16 | scratch 2*NWARPS*6 doubles |
//   coff nbuf * n_a bytes (padded to 16) | crop buffers nbuf * n_a * pitch doubles ]
// ---------------------------------------------------------------------------------------------
struct FitArgs {
    const double *stamps;
    int r, c, n_a;
    long long N;
    const double *x0;
    double g_tol;
    int max_iter;
    double *s, *g1, *g2;
    int *status, *iters;
    unsigned long long *counter;   // device work counter, zeroed before launch
    int nbuf, pitch, plane;
    double *f_out, *g_out;         // non-null: evaluate cost / gradient at x0 only (cost, g! of analyticLBFGS.jl)
    int profile;                   // 0: fGaussian (radialProfiles.jl:5-9), 1: fkolmogorov (:16-24)
    int sym;
};

// Issue the bulk copies of one star's crop (thread 0 only).  A column whose 16-byte-aligned superset would
// leave the batch buffer is skipped here and loaded with plain loads by load_edges().
__device__ __forceinline__ void stage_issue(const FitArgs &a, long long star, double *crop, unsigned char *coff,
                                            uint64_t *bar) {
    const int r0 = (a.r - a.n_a) / 2, c0 = (a.c - a.n_a) / 2;   // get_middle_nxn: div(rows-n,2)+1, 1-based
    const long long total = a.N * (long long)a.r * a.c;
    const long long base = star * (long long)a.r * a.c + r0 + (long long)c0 * a.r;
    const int mis0 = (int)(((uintptr_t)a.stamps >> 3) & 1);
    uint32_t bytes = 0;
    for (int v = 0; v < a.n_a; v++) {
        const long long ge = base + (long long)v * a.r;
        const int m = (int)((ge + mis0) & 1);
        const long long start = ge - m;
        const int len = (m + a.n_a + 1) & ~1;
        coff[v] = (unsigned char)m;
        if (start >= 0 && start + len <= total) {
            bulk_g2s(crop + (size_t)v * a.pitch, a.stamps + start, (uint32_t)len * 8u, bar);
            bytes += (uint32_t)len * 8u;
        }
    }
    mbar_arrive_expect_tx(bar, bytes);
}
// plain-load fallback for the (at most two) columns stage_issue skipped; the slot's master warp
__device__ __forceinline__ void load_edges(const FitArgs &a, long long star, double *crop) {
    const int r0 = (a.r - a.n_a) / 2, c0 = (a.c - a.n_a) / 2;
    const long long total = a.N * (long long)a.r * a.c;
    const long long base = star * (long long)a.r * a.c + r0 + (long long)c0 * a.r;
    const int mis0 = (int)(((uintptr_t)a.stamps >> 3) & 1);
    for (int v = 0; v < a.n_a; v += (a.n_a > 1 ? a.n_a - 1 : 1)) {   // only the first and last column can be edges
        const long long ge = base + (long long)v * a.r;
        const int m = (int)((ge + mis0) & 1);
        const long long start = ge - m;
        const int len = (m + a.n_a + 1) & ~1;
        if (!(start >= 0 && start + len <= total)) {
            for (int u = threadIdx.x & 31; u < a.n_a; u += 32) crop[(size_t)v * a.pitch + m + u] = a.stamps[ge + u];
        }
    }
}

// Shared memory (dynamic): [ NSLOT x (EvalShared | scratch 2*NWARPS*6 doubles) | mbarriers NSLOT x 8 B, padded to 16 |
//   coff NSLOT x stride | crops NSLOT x n_a*pitch doubles | profile plane n_a*pitch doubles (shared by the slots) |
//   exp table 64 x 16 doubles ]
template <int THREADS, int NSLOT, int MINB>
__global__ void __launch_bounds__(THREADS + 32 * NSLOT, MINB) fit_kernel(const FitArgs a) {
    unsigned char *p = smem_raw + NSLOT * slot_stride<THREADS>();
    uint64_t *bars = reinterpret_cast<uint64_t *>(p);
    unsigned char *coff_base = p + ((NSLOT * 8 + 15) & ~15);
    const int coff_stride = (a.n_a + 15) & ~15;
    double *crop_base = reinterpret_cast<double *>(coff_base + (size_t)NSLOT * coff_stride);
    const size_t crop_stride = (size_t)a.n_a * a.pitch;
    double *fplane = crop_base + (size_t)NSLOT * crop_stride;
    double *exp_tab = fplane + (size_t)a.plane * crop_stride;
    for (int i = threadIdx.x; i < EXP_TAB_N * EXP_TAB_REP; i += blockDim.x) exp_tab[i] = EXP_T64[i / EXP_TAB_REP];

    if (threadIdx.x == 0) {
        for (int k = 0; k < NSLOT; k++) mbar_init(&bars[k], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x < THREADS) {
        worker_loop<THREADS, NSLOT>(bars);
        return;
    }
    // ---- master warp of slot `slot`: its own star loop; every lane runs the scalar state machine redundantly ----
    const int slot = master_slot<THREADS>(), lane = threadIdx.x & 31;
    EvalShared &es = eval_shared<THREADS>(slot);
    double *crop = crop_base + slot * crop_stride;
    unsigned char *coff = coff_base + slot * coff_stride;
    const int r0 = (a.r - a.n_a) / 2, c0 = (a.c - a.n_a) / 2;
    const int mis0 = (int)(((uintptr_t)a.stamps >> 3) & 1);
    uint32_t phase = 0u;
    long long nx = 0;
    if (lane == 0) {
        nx = (long long)atomicAdd(a.counter, 1ULL);
        if (nx < a.N) stage_issue(a, nx, crop, coff, &bars[slot]);
    }
    for (;;) {
        const long long cur = __shfl_sync(0xffffffffu, nx, 0);
        if (cur >= a.N) break;
        if (lane == 0) {
            const long long base = cur * (long long)a.r * a.c + r0 + (long long)c0 * a.r;
            es.crop_s = smem_u32(crop); es.fbuf_s = smem_u32(fplane); es.tab_s = smem_u32(exp_tab);
            es.n_a = a.n_a; es.pitch = a.pitch; es.TV = THREADS / a.n_a;      // n_a <= THREADS (checked at launch)
            es.m0 = (int)((base + mis0) & 1); es.rodd = a.r & 1; es.plane = a.plane;
            es.have_last = 0; es.cmd = 0; es.newstar = 1; es.parity = (int)phase; es.profile = a.profile; es.sym = a.sym;
        }
        mbar_wait(&bars[slot], phase);
        phase ^= 1u;
        load_edges(a, cur, crop);
        __syncwarp();
        const double x0[3] = {a.x0[3 * cur + 0], a.x0[3 * cur + 1], a.x0[3 * cur + 2]};
        double pr[3];
        int iters = 0;
        if (a.f_out) {   // cost(x0), g!(G, x0) only
            const FG r = eval_fg<THREADS>(x0[0], x0[1], x0[2]);
            if (lane == 0) {
                a.f_out[cur] = r.f;
                a.g_out[3 * cur + 0] = r.g0; a.g_out[3 * cur + 1] = r.g1; a.g_out[3 * cur + 2] = r.g2;
            }
        }
        const int st = a.f_out ? 0 : lbfgs_fit<THREADS>(x0, a.g_tol, a.max_iter, pr, iters);
        if (lane == 0) {
            if (!a.f_out) {
                a.s[cur] = pr[0]; a.g1[cur] = pr[1]; a.g2[cur] = pr[2];
                if (a.status) a.status[cur] = st;
                if (a.iters) a.iters[cur] = iters;
            }
            // the workers left this slot's crop at the last evaluation's final barrier: the next star may land on it
            nx = (long long)atomicAdd(a.counter, 1ULL);
            if (nx < a.N) {
                fence_proxy_async();
                stage_issue(a, nx, crop, coff, &bars[slot]);
            }
        }
        __syncwarp();
    }
    if (lane == 0) es.cmd = CMD_EXIT;
    __syncwarp();
    bar_eval<THREADS>(slot);   // tells the workers this slot is finished
}

// ---- launch ---------------------------------------------------------------------------------
template <int THREADS>
static size_t fit_smem_bytes(int nslot, int n_a, int pitch, int plane) {
    const size_t coff_stride = (size_t)((n_a + 15) & ~15);
    return (size_t)nslot * slot_stride<THREADS>() + (size_t)((nslot * 8 + 15) & ~15) + nslot * coff_stride +
           (size_t)(nslot + plane) * n_a * pitch * sizeof(double) + EXP_TAB_BYTES;
}

template <int THREADS, int NSLOT, int MINB>
static int launch_fit_cfg(FitArgs a, cudaStream_t stream, bool probe_only, int *per_sm_out) {
    if (a.n_a > THREADS) return SHOPT_E_UNSUPPORTED;
    a.pitch = (a.n_a + 2 + 1) & ~1;
    auto kern = fit_kernel<THREADS, NSLOT, MINB>;
    // the profile plane is dropped (pass B recomputes exp) when crops + plane do not fit one SM
    a.plane = fit_smem_bytes<THREADS>(NSLOT, a.n_a, a.pitch, 1) <= 227 * 1024 ? 1 : 0;
    const size_t smem = fit_smem_bytes<THREADS>(NSLOT, a.n_a, a.pitch, a.plane);
    if (smem > 227 * 1024) return SHOPT_E_UNSUPPORTED;
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    SHOPT_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, THREADS + 32 * NSLOT, smem));
    if (per_sm_out) *per_sm_out = per_sm * (a.plane ? 2 : 1);    // a configuration that keeps the plane is preferred
    if (per_sm < 1) return SHOPT_E_UNSUPPORTED;
    if (probe_only) return SHOPT_OK;
    long long grid = (long long)sm_count() * per_sm;
    const long long want = (a.N + NSLOT - 1) / NSLOT;
    if (grid > want) grid = want;
    SHOPT_CUDA_TRY(cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), stream));
    kern<<<(unsigned)grid, THREADS + 32 * NSLOT, smem, stream>>>(a);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
// two stars per CTA when two crops and the plane fit, else one
template <int THREADS, int MINB2, int MINB1>
static int launch_fit(FitArgs a, cudaStream_t stream) {
    static const int force = getenv("SHOPT_FIT_NSLOT") ? atoi(getenv("SHOPT_FIT_NSLOT")) : 0;   // measurement switch
    int p2 = 0, p1 = 0;
    const int r2 = (force == 1) ? SHOPT_E_UNSUPPORTED : launch_fit_cfg<THREADS, 2, MINB2>(a, stream, true, &p2);
    const int r1 = (force == 2) ? SHOPT_E_UNSUPPORTED : launch_fit_cfg<THREADS, 1, MINB1>(a, stream, true, &p1);
    if (r2 == SHOPT_OK && (r1 != SHOPT_OK || 2 * p2 >= p1)) return launch_fit_cfg<THREADS, 2, MINB2>(a, stream, false, nullptr);
    if (r1 == SHOPT_OK) return launch_fit_cfg<THREADS, 1, MINB1>(a, stream, false, nullptr);
    return SHOPT_E_UNSUPPORTED;
}

int fit_analytic_launch(const double *d_stamps, int r, int c, long long N, int n_a, const double *d_x0, double g_tol,
                        int max_iter, double *d_s, double *d_g1, double *d_g2, int *d_status, int *d_iters,
                        unsigned long long *d_counter, cudaStream_t stream, double *d_f, double *d_g, int profile) {
    if (!d_stamps || !d_x0 || !d_counter || profile < 0 || profile > 1) return SHOPT_E_BADARG;
    if (d_f ? !d_g : (!d_s || !d_g1 || !d_g2)) return SHOPT_E_BADARG;
    if (r < 1 || c < 1 || n_a < 1 || n_a > r || n_a > c || N < 0 || max_iter < 0) return SHOPT_E_BADARG;
    if (N == 0) return SHOPT_OK;
    FitArgs a;
    a.stamps = d_stamps; a.r = r; a.c = c; a.n_a = n_a; a.N = N; a.x0 = d_x0; a.g_tol = g_tol; a.max_iter = max_iter;
    a.s = d_s; a.g1 = d_g1; a.g2 = d_g2; a.status = d_status; a.iters = d_iters; a.counter = d_counter;
    a.nbuf = 2; a.pitch = 0; a.plane = 1; a.f_out = d_f; a.g_out = d_g; a.profile = profile;
    static const int sym = getenv("SHOPT_FIT_SYM") ? atoi(getenv("SHOPT_FIT_SYM")) : 1;   // measurement switch
    a.sym = sym;
    if (n_a <= 32) return launch_fit<128, 4, 6>(a, stream);
    if (n_a <= 80) return launch_fit<256, 2, 3>(a, stream);
    if (n_a <= 160) return launch_fit<512, 1, 1>(a, stream);
    return SHOPT_E_UNSUPPORTED;
}

}  // namespace shopt

// pca_power.cu -- stage 1 of the PCA pixel-grid fit (shopt/shopt.jl:281-314; pca_image, shopt/pca.jl:1-18): the stars
// whose reconstruction has ONE component, which on a catalogue of well-sampled stars is every star.
//
// pca_kernel (pixelgrid.cu) keeps the centred stamp Z in shared memory (137 KB at 131 x 131), so one star at a time lives
// on an SM and every phase of it -- load, mean, power iteration, reconstruction -- waits on its own dependent chain:
// FP64 pipe 10 %, issue slots 31 %, 86k cycles per star.  For a one-component star nothing needs Z resident: the power
// iteration is a handful of passes over the stamp, and after the first pass the stamp sits in L2.  This kernel therefore
// holds NO copy of the stamp: z_ij = nanToZero(x_ij) - mean_i is rebuilt from global memory (L2) in every pass, a CTA is
// four warps with a few KB of shared memory, and six or more stars are in flight per SM, which hides the latency that
// bounds the one-star-per-SM kernel.  Lane l owns rows l, l + 32, ... (its slices of q, the mean and Z t stay in
// registers), warp w owns columns w, w + 4, ...: one pass computes  t = Z'q  and  Z t  together (a column is loaded once:
// dot with q, butterfly, then multiplied into the row accumulators).
//
// Same mathematics and the same A-POSTERIORI acceptance as power_pca in pixelgrid.cu: k = 1 is taken only when
// theta = |Z'q|^2 (a lower bound of sigma_1^2) already reaches pratio of |Z|_F^2 with the 1e-11 guard band of
// choose_pcadim's comparison and |Z Z'q - theta q| <= 1e-13 |Z|_F^2.  Every other star -- more components, masked-pixel
// outliers, non-finite pixels, non-finite factors -- is appended to a work list and served by pca_kernel afterwards, which
// also owns the failure semantics (shopt.jl:305-309).  Output: smooth(mean 1' + u t') from the 1-D smoothed factors
// (dataPreprocessing.jl:112-139 is separable with a separable normalisation), or the reconstruction divided by its sum
// when unity_sum (the reference's one-argument smooth() call raises: shopt.jl:292-295).
#include <cstdlib>
#include <type_traits>

#include "api_ctx.h"
#include "common.cuh"
#include "kernels.h"

namespace shopt {

constexpr int PW_NW = 4, PW_THREADS = PW_NW * 32;

__device__ __forceinline__ double pw_ntz(double x) { return isnan(x) ? 0.001 : x; }   // masks.jl:43

// D > 0: the columns of the next D - 1 steps of a warp's walk over the stamp are in flight all the time: every lane copies
// ITS OWN rows of the two columns of a step into a private slot ring in shared memory (8-byte cp.async; a lane reads back
// only what it copied, so cp.async.wait_group is the only synchronisation).  Every pass walks the same columns in the same
// order, so the ring runs on across the passes of a star: while the CTA reduces and decides after a pass, the first
// columns of the next pass are already on their way (ncu on the D = 0 kernel: 41 % of the warp samples on the long
// scoreboard).  The values and the arithmetic are those of the D = 0 kernel: results are bit-identical.
// shared memory (doubles): mu r | q r | qn r | part NW*r | part2 NW*r | t c | Ys c | red 2*NW*3 | ring NW*D*2*RT*32
template <int RT, int D>
__global__ void __launch_bounds__(PW_THREADS, RT <= 5 ? (D > 0 ? 4 : 5) : 3) pca_power_kernel(const PcaPowerArgs a) {
    extern __shared__ __align__(16) double pw_smem[];
    __shared__ long long s_star;
    const int r = a.r, c = a.c;
    const long long rc = (long long)r * c;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    double *mu = pw_smem, *q = mu + r, *qn = q + r, *part = qn + r, *part2 = part + PW_NW * r, *t = part2 + PW_NW * r, *Ys = t + c,
           *red = Ys + c;
    int buf = 0;
    const double den = (double)(c - 1);
    // ring: slot s of this warp holds step columns (j, j + NW) as [s][which][k][lane]
    double *const ringw = red + 2 * PW_NW * 3 + (size_t)warp * (D * 2 * RT * 32) + lane;
    const int niter = (c - warp + 2 * PW_NW - 1) / (2 * PW_NW);   // steps of this warp per pass (c >= 2 NW: at least one)
    for (;;) {
        __syncthreads();                                         // the previous star's shared data is dead
        if (tid == 0) s_star = (long long)atomicAdd(a.counter, 1ULL);
        __syncthreads();
        const long long star = s_star;
        if (star >= a.N) break;
        const double *X = a.stamps + star * rc;
        double *out = a.pgf + star * rc;
        bool defer = false;
        int pn = 0, ps = 0, cs = 0;                              // ring: next step to prefetch, its slot, the slot to read next
        auto issue = [&]() {
            if constexpr (D > 0) {
                const int j = warp + 2 * PW_NW * pn;
                const double *c1 = X + lane + (size_t)j * r;
                double *dst = ringw + ps * (2 * RT * 32);
                const bool two = j + PW_NW < c;
#pragma unroll
                for (int k = 0; k < RT; k++)
                    if (k < RT - 1 || lane + 32 * k < r) {
                        cp_async8(dst + k * 32, c1 + 32 * k);
                        if (two) cp_async8(dst + (RT + k) * 32, c1 + (size_t)PW_NW * r + 32 * k);
                    }
                cp_async_commit();
                if (++pn == niter) pn = 0;
                if (++ps == D) ps = 0;
            }
        };
        auto landed = [&]() -> const double * {                  // one more step out, then the oldest one has arrived
            issue();
            if constexpr (D > 0) cp_async_wait<D - 1>();
            const double *src = ringw + cs * (2 * RT * 32);
            if (++cs == D) cs = 0;
            return src;
        };
        if constexpr (D > 0)
            for (int d = 0; d < D - 1; d++) issue();
        // ---- pass 1 (HBM): row sums -> mean over the columns; masked pixels, non-finite pixels; AND a first power step.
        //      With q0 = the (uncentred) central column, normalised, and a = X'q0:  Z Z'q0 = X a - mean (1'a)  for
        //      Z = X - mean 1' (the terms in mean'q0 cancel), so the first application of Z Z' needs no mean and rides
        //      along with the pass that computes it.  It only supplies the start vector of the iteration below (whose
        //      passes use the centred pixels and decide acceptance a posteriori), and saves one pass over the stamp. ----
        double nrep = 0.0;
        bool hasnan = false;
        {
            const double *xl = X + lane;
            const double *xc = xl + (size_t)(c / 2) * r;
            double q0[RT], v0[3] = {0.0, 0.0, 0.0};
#pragma unroll
            for (int k = 0; k < RT; k++) {
                q0[k] = (k < RT - 1 || lane + 32 * k < r) ? pw_ntz(__ldg(xc + 32 * k)) : 0.0;
                if (warp == 0) v0[0] = fma(q0[k], q0[k], v0[0]);
            }
            cta_sum<3, PW_NW>(v0, red, buf);
            const double inv0 = (v0[0] > 0.0 && isfinite(v0[0])) ? 1.0 / sqrt(v0[0]) : 0.0;   // 0: the later checks defer the star
#pragma unroll
            for (int k = 0; k < RT; k++) q0[k] *= inv0;
            double rs[RT], xa[RT];
#pragma unroll
            for (int k = 0; k < RT; k++) { rs[k] = 0.0; xa[k] = 0.0; }
            int nrepi = 0;
            double suma = 0.0;
            const bool hi16 = lane & 16;
            for (int j = warp; j < c; j += 2 * PW_NW) {
                const int j2 = j + PW_NW;
                const bool two = j2 < c;
                const double *c1 = xl + (size_t)j * r, *c2 = xl + (size_t)(two ? j2 : j) * r;
                double x1[RT], x2[RT], d1 = 0.0, d2 = 0.0;
                const double *src = nullptr;
                if constexpr (D > 0) src = landed();
#pragma unroll
                for (int k = 0; k < RT; k++) {
                    x1[k] = 0.0; x2[k] = 0.0;
                    if (k < RT - 1 || lane + 32 * k < r) {
                        double y1, y2;
                        if constexpr (D > 0) { y1 = src[k * 32]; y2 = two ? src[(RT + k) * 32] : 0.0; }
                        else { y1 = __ldg(c1 + 32 * k); y2 = __ldg(c2 + 32 * k); }
                        const bool n1 = y1 != y1, n2 = two && (y2 != y2);
                        nrepi += n1; nrepi += n2;
                        x1[k] = n1 ? 0.001 : y1;
                        x2[k] = two ? (n2 ? 0.001 : y2) : 0.0;
                    }
                    rs[k] += x1[k] + x2[k];
                    d1 = fma(x1[k], q0[k], d1);
                    d2 = fma(x2[k], q0[k], d2);
                }
                double sx = (hi16 ? d2 : d1) + __shfl_xor_sync(0xffffffffu, hi16 ? d1 : d2, 16);
                sx += __shfl_xor_sync(0xffffffffu, sx, 8);
                sx += __shfl_xor_sync(0xffffffffu, sx, 4);
                sx += __shfl_xor_sync(0xffffffffu, sx, 2);
                sx += __shfl_xor_sync(0xffffffffu, sx, 1);
                d1 = __shfl_sync(0xffffffffu, sx, 0);
                d2 = __shfl_sync(0xffffffffu, sx, 16);
#pragma unroll
                for (int k = 0; k < RT; k++) xa[k] = fma(x2[k], d2, fma(x1[k], d1, xa[k]));
                if (lane == 0) suma += d1 + d2;
            }
            double tot = 0.0;                                    // an Inf anywhere makes this row sum non-finite
#pragma unroll
            for (int k = 0; k < RT; k++) {
                const int i = lane + 32 * k;
                if (i < r) { part[warp * r + i] = rs[k]; part2[warp * r + i] = xa[k]; }
                tot += rs[k];
            }
            nrep = (double)nrepi;
            hasnan = __syncthreads_or(nrepi) != 0;
            if (__syncthreads_or(!isfinite(tot))) defer = true;  // LAPACK raises on Inf: pca_kernel writes the failure
            double v1[3] = {suma, 0.0, 0.0};
            cta_sum<3, PW_NW>(v1, red, buf);
            if (!defer) {
                double v2[3] = {0.0, 0.0, 0.0};
                for (int i = tid; i < r; i += PW_THREADS) {
                    double sr = part[i], sa = part2[i];
#pragma unroll
                    for (int w = 1; w < PW_NW; w++) { sr += part[w * r + i]; sa += part2[w * r + i]; }
                    const double m = sr / c;
                    mu[i] = m;
                    const double z = sa - m * v1[0];             // (Z Z'q0)_i
                    qn[i] = z;
                    v2[0] = fma(z, z, v2[0]);
                }
                cta_sum<3, PW_NW>(v2, red, buf);
                if (!(v2[0] > 0.0) || !isfinite(v2[0])) defer = true;
                else {
                    const double inv = 1.0 / sqrt(v2[0]);
                    for (int i = tid; i < r; i += PW_THREADS) q[i] = qn[i] * inv;
                }
                __syncthreads();
            }
        }
        // ---- power iteration: one pass per step computes t = Z'q and Z t together ----
        double theta = 0.0, fro = 0.0;
        bool accepted = false;
        double mur[RT];
#pragma unroll
        for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; mur[k] = (!defer && i < r) ? mu[i] : 0.0; }
        for (int it = 1; it <= 8 && !defer && !accepted; it++) {
            double qr[RT], qa[RT];
#pragma unroll
            for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; qr[k] = i < r ? q[i] : 0.0; qa[k] = 0.0; }
            double th = 0.0, frp = 0.0;
            const double *xl = X + lane;
            const bool hi16 = lane & 16;
            auto pass = [&](auto NANS) {
                constexpr bool nans = decltype(NANS)::value;
                for (int j = warp; j < c; j += 2 * PW_NW) {
                    const int j2 = j + PW_NW;
                    const bool two = j2 < c;
                    const double *c1 = xl + (size_t)j * r, *c2 = xl + (size_t)(two ? j2 : j) * r;
                    double z1[RT], z2[RT], d1 = 0.0, d2 = 0.0;
                    const double *src = nullptr;
                    if constexpr (D > 0) src = landed();
#pragma unroll
                    for (int k = 0; k < RT; k++) {
                        z1[k] = 0.0; z2[k] = 0.0;
                        if (k < RT - 1 || lane + 32 * k < r) {
                            double x1, x2;
                            if constexpr (D > 0) { x1 = src[k * 32]; x2 = two ? src[(RT + k) * 32] : 0.0; }
                            else { x1 = __ldg(c1 + 32 * k); x2 = __ldg(c2 + 32 * k); }
                            if (nans) { x1 = pw_ntz(x1); x2 = pw_ntz(x2); }
                            z1[k] = x1 - mur[k];
                            z2[k] = two ? x2 - mur[k] : 0.0;
                        }
                        d1 = fma(z1[k], qr[k], d1);
                        d2 = fma(z2[k], qr[k], d2);
                        if (it == 1) { frp = fma(z1[k], z1[k], frp); frp = fma(z2[k], z2[k], frp); }
                    }
                    // two sums for the price of one butterfly: the halves of the warp take one each
                    double sx = (hi16 ? d2 : d1) + __shfl_xor_sync(0xffffffffu, hi16 ? d1 : d2, 16);
                    sx += __shfl_xor_sync(0xffffffffu, sx, 8);
                    sx += __shfl_xor_sync(0xffffffffu, sx, 4);
                    sx += __shfl_xor_sync(0xffffffffu, sx, 2);
                    sx += __shfl_xor_sync(0xffffffffu, sx, 1);
                    d1 = __shfl_sync(0xffffffffu, sx, 0);
                    d2 = __shfl_sync(0xffffffffu, sx, 16);
#pragma unroll
                    for (int k = 0; k < RT; k++) qa[k] = fma(z2[k], d2, fma(z1[k], d1, qa[k]));
                    if (lane == 0) {
                        t[j] = d1; th = fma(d1, d1, th);
                        if (two) { t[j2] = d2; th = fma(d2, d2, th); }
                    }
                }
            };
            if (hasnan) pass(std::true_type{}); else pass(std::false_type{});
#pragma unroll
            for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; if (i < r) part[warp * r + i] = qa[k]; }
            double v3[3] = {th, frp, (it == 1) ? nrep : 0.0};
            cta_sum<3, PW_NW>(v3, red, buf);                     // its barrier also publishes part and t
            theta = v3[0];
            if (it == 1) {
                fro = v3[1];
                // masked pixels (0.001 each) that alone exceed half of the variance k components may leave out: many
                // components, pca_kernel's direct path (the same routing rule as in pca_kernel)
                if (!isfinite(fro) || v3[2] * 1e-6 > 0.5 * (1.0 - a.pratio) * fro) { defer = true; break; }
            }
            const double thres = (fro / den) * a.pratio, tol = 1e-13 * fro;
            double v2[3] = {0.0, 0.0, 0.0};
            for (int i = tid; i < r; i += PW_THREADS) {
                double x = part[i];
#pragma unroll
                for (int w = 1; w < PW_NW; w++) x += part[w * r + i];
                qn[i] = x;
                const double d = x - theta * q[i];
                v2[0] = fma(d, d, v2[0]);
                v2[1] = fma(x, x, v2[1]);
            }
            cta_sum<3, PW_NW>(v2, red, buf);
            const double acc1 = theta / den;
            if (!(acc1 < thres) && !(fabs(acc1 - thres) <= 1e-11 * thres) && v2[0] <= tol * tol && theta > 0.0) {
                accepted = true;
                break;
            }
            // a Rayleigh quotient well below the threshold after two steps will not reach it: not a one-component star
            if (!(v2[1] > 0.0) || !isfinite(v2[1]) || (it >= 2 && acc1 < 0.97 * thres)) { defer = true; break; }
            const double inv = 1.0 / sqrt(v2[1]);
            __syncthreads();                                     // everyone has read q (d above) before it changes
            for (int i = tid; i < r; i += PW_THREADS) q[i] = qn[i] * inv;
            __syncthreads();
        }
        if (!accepted) defer = true;
        // ---- output (k = 1): u = q, y = t = Z'u ----
        const bool smoothing = !a.unity_sum && a.lanczos_a > 0;
        if (!defer && smoothing) {
            // smooth(mean 1' + u y') = M 1' + U' Y' with the 1-D smoothed factors (taps inside the image, renormalised)
            const int la = a.lanczos_a, kw = 2 * la + 1;
            int nf = 0;
            double *Ms = qn, *Us = part;                         // both free now
            for (int e = tid; e < 2 * r; e += PW_THREADS) {
                const int which = e / r, i = e - which * r;
                const double *src = which ? q : mu;
                double acc = 0.0, A = 0.0;
                for (int k = 0; k < kw; k++) {
                    const int ii = i + k - la;
                    if (ii >= 0 && ii < r) { acc = fma(a.kz[k], src[ii], acc); A += a.kz[k]; }
                }
                const double val = acc / A;
                if (!isfinite(val)) nf = 1;
                (which ? Us : Ms)[i] = val;
            }
            for (int j = tid; j < c; j += PW_THREADS) {
                double acc = 0.0, B = 0.0;
                for (int k = 0; k < kw; k++) {
                    const int jj = j + k - la;
                    if (jj >= 0 && jj < c) { acc = fma(a.kz[k], t[jj], acc); B += a.kz[k]; }
                }
                const double val = acc / B;
                if (!isfinite(val)) nf = 1;
                Ys[j] = val;
            }
            nf = __syncthreads_or(nf);                           // also publishes the factors
            if (nf) defer = true;
            else {
                double mr[RT], ur[RT];
#pragma unroll
                for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; mr[k] = i < r ? Ms[i] : 0.0; ur[k] = i < r ? Us[i] : 0.0; }
                for (int j = warp; j < c; j += PW_NW) {
                    const double yj = Ys[j];
                    double *o = out + (size_t)j * r;
#pragma unroll
                    for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; if (i < r) __stcs(o + i, fma(ur[k], yj, mr[k])); }
                }
            }
        } else if (!defer) {
            // the reconstruction itself, divided by its sum when unity_sum (shopt.jl:295: no smoothing in that branch)
            double ur[RT];
#pragma unroll
            for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; ur[k] = i < r ? q[i] : 0.0; }
            double scale = 1.0;
            if (a.unity_sum) {
                double v1[3] = {0.0, 0.0, 0.0};
                for (int j = warp; j < c; j += PW_NW) {
                    const double yj = t[j];
#pragma unroll
                    for (int k = 0; k < RT; k++) { const int i = lane + 32 * k; if (i < r) v1[0] += fma(ur[k], yj, mur[k]); }
                }
                cta_sum<3, PW_NW>(v1, red, buf);
                scale = v1[0];
            }
            for (int j = warp; j < c; j += PW_NW) {
                const double yj = t[j];
                double *o = out + (size_t)j * r;
#pragma unroll
                for (int k = 0; k < RT; k++) {
                    const int i = lane + 32 * k;
                    if (i < r) { const double val = fma(ur[k], yj, mur[k]); __stcs(o + i, a.unity_sum ? val / scale : val); }
                }
            }
        }
        if constexpr (D > 0) cp_async_wait<0>();                 // the steps prefetched for a pass that did not happen
        if (tid == 0) {
            if (defer) {
                a.defer_list[atomicAdd(a.defer_count, 1ULL)] = star;
            } else {
                if (a.k_out) a.k_out[star] = 1;
                if (a.status) a.status[star] = SHOPT_PCA_OK;
            }
        }
    }
}

template <int RT, int D>
static int pca_power_launch_d(const PcaPowerArgs &a, cudaStream_t st) {
    const size_t smem = ((size_t)(3 + 2 * PW_NW) * a.r + (size_t)2 * a.c + 2 * PW_NW * 3 + (size_t)PW_NW * D * 2 * RT * 32) * sizeof(double);
    auto kern = pca_power_kernel<RT, D>;
    if (smem > 48 * 1024) SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    SHOPT_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, PW_THREADS, smem));
    if (per_sm < 1) return SHOPT_E_UNSUPPORTED;
    // Every pass after the first should find the stamp in L2, so the stamps in flight (one per resident CTA) are held to
    // ~96 MB of the 126 MB L2 (131 x 131: 4 CTAs per SM measured 15.0 ms per 100k stars, 5: 16.3, 3: 17.8, 2: 24.8).
    // Stamps too large for that (301 x 301: 0.9 per SM) stream from HBM in every pass anyway: occupancy decides.
    static const int cap_env = getenv("SHOPT_PCA_LIGHT_CTAS") ? atoi(getenv("SHOPT_PCA_LIGHT_CTAS")) : 0;   // measurement switch
    int cap = cap_env;
    if (cap <= 0) {
        const double fit = 96e6 / ((double)sm_count() * (double)a.r * (double)a.c * 8.0);
        cap = fit >= 2.0 ? (int)fit : 0;
    }
    if (cap > 0 && per_sm > cap) per_sm = cap;
    long long grid = (long long)sm_count() * per_sm;
    if (grid > a.N) grid = a.N;
    SHOPT_CUDA_TRY(cudaMemsetAsync(a.counter, 0, 2 * sizeof(unsigned long long), st));   // work counter, defer count
    kern<<<(unsigned)grid, PW_THREADS, smem, st>>>(a);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

// SHOPT_PCA_RING=0 / 1 forces the plain-load kernel (D = 0) / the ring.  Measured (B200, per 100k stars of 131 x 131):
// 13.14 ms plain, 13.24 ms ring -- at four stamps per SM living in L2 the long-scoreboard samples overlap with the
// butterfly chains and the CTA-wide reductions, and the ring's extra LDS / LDGSTS cost what the prefetch wins; per 6 250
// stars of 301 x 301 (every pass streams from HBM): 4.52 -> 4.28 ms.  Default: the ring for stamps above 160 rows.
template <int RT>
static int pca_power_launch_t(const PcaPowerArgs &a, cudaStream_t st) {
    static const int ring = getenv("SHOPT_PCA_RING") ? atoi(getenv("SHOPT_PCA_RING")) : -1;
    constexpr int DD = RT <= 5 ? 4 : 3;
    if (ring > 0 || (ring < 0 && RT > 5)) return pca_power_launch_d<RT, DD>(a, st);
    return pca_power_launch_d<RT, 0>(a, st);
}

int pca_power_launch(const PcaPowerArgs &a, cudaStream_t st) {
    if (!a.stamps || !a.pgf || !a.counter || !a.defer_list || !a.defer_count || a.N <= 0 || a.lanczos_a > 7) return SHOPT_E_BADARG;
    if (a.defer_count != a.counter + 1) return SHOPT_E_BADARG;   // zeroed together
    const int rt = (a.r + 31) / 32;
    switch (rt) {
        case 1: return pca_power_launch_t<1>(a, st);
        case 2: return pca_power_launch_t<2>(a, st);
        case 3: return pca_power_launch_t<3>(a, st);
        case 4: return pca_power_launch_t<4>(a, st);
        case 5: return pca_power_launch_t<5>(a, st);
        case 6: case 7: return pca_power_launch_t<7>(a, st);
        case 8: case 9: case 10: return pca_power_launch_t<10>(a, st);
        default: return SHOPT_E_UNSUPPORTED;
    }
}

}  // namespace shopt

// pixelgrid.cu -- per-star pixel-grid fit in "PCA" mode, and the standalone smooth().
//
// Replaces the loop shopt/shopt.jl:281-314 of the reference:
//     data = nanToZero(star)                       shopt/masks.jl:38-50      (NaN -> 0.001)
//     pca_image(data, PCAterms)                    shopt/pca.jl:1-18         (MultivariateStats fit/transform/reconstruct,
//                                                                             rows = variables, columns = observations)
//     smooth(., lanczos)  or  ./ sum(.)            shopt/dataPreprocessing.jl:94-139, shopt.jl:292-302
// MultivariateStats 0.10.2 is not vendored; the algorithm is the one recorded in SURVEY.md A.3: mean over the columns,
// SVD of the centred matrix Z, variances sigma_i^2/(c-1), k = first count reaching pratio of the total (capped by
// ncomp and min(r,c)), output mean + U_k U_k' Z.
//
// Design (B200): one persistent CTA per star.  The centred stamp W = Z (r x c, column-major) lives in shared memory
// (137 KB at 131 x 131) and is diagonalised in place by ONE-SIDED (Hestenes) JACOBI rotations: a sweep is c-1 (or c)
// steps of a round-robin tournament, each step rotating floor(c/2) disjoint column pairs, one warp per pair (the three
// dot products and the rotation keep the two columns in registers).  On convergence the columns are sigma_j u_j, which
// is all PCA needs: variances are the squared column norms, and the projection  Y = U_k' Z  re-reads the stamp from
// global memory (L2-hot), so no second r x c buffer and no V accumulation is needed.  The reconstruction is written to
// the output, pulled back into the (now free) shared buffer and smoothed from there.  Stamps too large for shared
// memory (301 x 301) run the same code with W in a per-CTA global scratch block that stays L2-resident.
// The stage is FP64-compute-bound (O(r c^2) flops per sweep on 8 r c bytes); see DESIGN.md for the roofline.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "pca_eig.cuh"

namespace shopt {

struct PcaArgs {
    const double *stamps;
    int r, c;
    long long N;
    int ncomp;
    double pratio;
    int lanczos_a, unity_sum;
    double *pgf;
    int *k_out, *status;
    unsigned long long *counter;
    double *wglobal;      // non-null: per-CTA scratch of wstride doubles replaces the shared-memory W / Y buffers
    size_t wstride;
    int kmax;             // min(ncomp, r, c)
    int fast;             // try the block subspace iteration first (0: always the full Jacobi SVD)
    int debug;            // status |= (subspace iterations used by the fast path; 100 direct path, 200 Jacobi) << 8
    double *eigscratch;   // non-null: per-CTA global scratch (eigstride doubles: Gram matrix + reflectors) of the direct path
    size_t eigstride;
    const long long *list;              // non-null: the stars to process are list[0 .. *nlist) (what pca_power_kernel
    const unsigned long long *nlist;    //   left over); null: stars 0 .. N-1
    double kern[15 * 15]; // normalised lanczos kernel (2a+1)^2, a <= 7, column-major like the reference's
    double kz[15];        // its 1-D factor: kern = kz (x) kz / sum
};

// SHOPT_PCA_DEBUG bit 16: cycles per phase of pca_kernel, summed over the stars (thread 0 of every CTA), printed by pca_launch
static __device__ unsigned long long g_pca_phase[16];   // one copy per translation unit (see SHOPT_PCA_RT below)

__device__ __forceinline__ double nan_to_zero(double x) { return isnan(x) ? 0.001 : x; }   // masks.jl:43

template <int NW>
__device__ __forceinline__ double block_sum1(double v, double *scratch, int &buf) {
    double a[1] = {v};
    cta_sum<1, NW>(a, scratch, buf);
    return a[0];
}

// ---------------------------------------------------------------------------------------------
// One-sided (Hestenes) Jacobi on the columns of A (nrows x ncols, leading dimension lda, shared or global memory):
// on return the columns are mutually orthogonal (A <- A V).  B, when given, is co-rotated (B <- B V).  A sweep is a
// round-robin tournament: npos-1 steps of npos/2 disjoint column pairs, one warp per pair.  Columns whose squared norm
// is <= tiny are left alone.  Must be entered by every thread of the CTA with A (and B) visible; returns the
// "still rotating" flag (non-zero when the sweep cap was hit).
// ---------------------------------------------------------------------------------------------
template <int NW, int RT>
__device__ int jacobi_columns(double *A, int nrows, int ncols, int lda, double *B, int nrowsB, int ldb, double tiny,
                              int max_sweeps, int &sweeps_out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int npos = (ncols & 1) ? ncols + 1 : ncols;     // one dummy position when ncols is odd
    const int nm1 = npos - 1, npairs = npos / 2;
    int sweeps = 0, rotated = 1;
    while (rotated && sweeps < max_sweeps) {
        rotated = 0;
        for (int step = 0; step < nm1; step++) {
            for (int pr = warp; pr < npairs; pr += NW) {
                int p, q;
                if (pr == 0) { p = nm1; q = step; }
                else { p = step + pr; if (p >= nm1) p -= nm1; q = step - pr; if (q < 0) q += nm1; }
                if (p > q) { const int t = p; p = q; q = t; }
                if (q >= ncols) continue;                   // the dummy position
                double *cp = A + (size_t)p * lda, *cq = A + (size_t)q * lda;
                double xa[RT], xb[RT];
                double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
                for (int t = 0; t < RT; t++) {
                    const int i = lane + 32 * t;
                    xa[t] = (i < nrows) ? cp[i] : 0.0;
                    xb[t] = (i < nrows) ? cq[i] : 0.0;
                    al = fma(xa[t], xa[t], al);
                    be = fma(xb[t], xb[t], be);
                    ga = fma(xa[t], xb[t], ga);
                }
                {   // three warp sums by one halving tree (6 shuffles instead of 15), then broadcast
                    const bool b4 = lane & 16, b3 = lane & 8;
                    const double k0 = (b4 ? ga : al) + __shfl_xor_sync(0xffffffffu, b4 ? al : ga, 16);
                    const double k1 = (b4 ? 0.0 : be) + __shfl_xor_sync(0xffffffffu, b4 ? be : 0.0, 16);
                    double k = (b3 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, b3 ? k0 : k1, 8);
                    k += __shfl_xor_sync(0xffffffffu, k, 4);
                    k += __shfl_xor_sync(0xffffffffu, k, 2);
                    k += __shfl_xor_sync(0xffffffffu, k, 1);
                    al = __shfl_sync(0xffffffffu, k, 0);      // lane 0: al, lane 8: be, lane 16: ga
                    be = __shfl_sync(0xffffffffu, k, 8);
                    ga = __shfl_sync(0xffffffffu, k, 16);
                }
                if (al <= tiny || be <= tiny) continue;
                if (ga * ga <= (DBL_EPSILON * DBL_EPSILON) * (al * be)) continue;
                rotated = 1;
                // tan of the rotation angle from  t = h sgn(d) / (|d| + sqrt(d^2 + h^2)),  d = be - al, h = 2 ga; only
                // cs^2 + sn^2 = 1 needs full accuracy (cs = rsqrt(1 + t^2), sn = t cs), so no IEEE division or square root
                const double dd = be - al, hh = 2.0 * ga;
                const double ww = fma(dd, dd, hh * hh);
                const double den = fabs(dd) + ww * rsqrt(ww);
                const double ri = rsqrt(den);
                double tt = copysign(hh, hh * dd) * (ri * ri);
                if (dd == 0.0) tt = copysign(1.0, ga);
                if (!isfinite(tt)) continue;
                const double cs = rsqrt(fma(tt, tt, 1.0)), sn = cs * tt;
#pragma unroll
                for (int t = 0; t < RT; t++) {
                    const int i = lane + 32 * t;
                    if (i < nrows) {
                        cp[i] = cs * xa[t] - sn * xb[t];
                        cq[i] = sn * xa[t] + cs * xb[t];
                    }
                }
                if (B) {
                    double *bp = B + (size_t)p * ldb, *bq = B + (size_t)q * ldb;
                    for (int i = lane; i < nrowsB; i += 32) {
                        const double x = bp[i], y = bq[i];
                        bp[i] = cs * x - sn * y;
                        bq[i] = sn * x + cs * y;
                    }
                }
            }
            __syncthreads();
        }
        rotated = __syncthreads_or(rotated);
        sweeps++;
    }
    sweeps_out = sweeps;
    return rotated;
}

// ---------------------------------------------------------------------------------------------
// Fast path: block subspace iteration for the few leading principal components.
//
// Star stamps are a handful of strong components over a noise floor, and pca_image keeps only the k components that
// reach pratio of the variance (k = 1..3 for the benchmark catalogues), so diagonalising all c columns is wasted work.
// The fast path iterates an FB = 8 column block  Q <- orth(Z Z' Q)  (two skinny products with Z in shared memory, a
// Cholesky-QR applied twice), then Rayleigh-Ritz on the block (one-sided Jacobi on the 8 columns of T = Z'Q, Q
// co-rotated) and ACCEPTS only when the answer is provably the one the full SVD gives:
//   * the Ritz values theta_1..theta_8 reach pratio of |Z|_F^2 within k <= 6 components, no comparison of the
//     cumulative variance against the threshold closer than 1e-11 relative;
//   * everything outside the block sums to E = |Z|_F^2 - sum(theta) <= theta_k / 4, so no component that was missed can
//     outrank the k kept ones (Ritz values are lower bounds of the eigenvalues they approximate);
//   * the residuals |Z Z' u_a - theta_a u_a| <= 1e-13 theta_1 for a <= k.
// Anything else (noise-dominated stamps, k > 6, slow convergence) falls through to the full Jacobi SVD below with Z
// untouched.  Buffers are column-major: Q, Qn (r x 8, leading dimension ldq), T (c x 8, leading dimension c).
// ---------------------------------------------------------------------------------------------
constexpr int FB = 8;
constexpr int FB_KMAX = 6;

// T = Z' Q : lanes = (column jj = lane & 7, row-part = lane >> 3); eight columns of Z per warp
template <int NW>
__device__ __forceinline__ void mul_zt_q(const double *Z, int r, int c, const double *Q, int ldq, double *T) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, jj = lane & 7, part = lane >> 3;
    for (int j0 = warp * 8; j0 < c; j0 += NW * 8) {
        const int j = j0 + jj;
        const bool live = j < c;
        const double *zc = Z + (size_t)(live ? j : 0) * r;
        double acc[FB];
#pragma unroll
        for (int b = 0; b < FB; b++) acc[b] = 0.0;
        for (int i = part; i < r; i += 4) {
            const double z = live ? zc[i] : 0.0;
#pragma unroll
            for (int b = 0; b < FB; b++) acc[b] = fma(z, Q[b * ldq + i], acc[b]);
        }
#pragma unroll
        for (int b = 0; b < FB; b++) {
            acc[b] += __shfl_xor_sync(0xffffffffu, acc[b], 8);
            acc[b] += __shfl_xor_sync(0xffffffffu, acc[b], 16);
        }
        if (part == 0 && live) {
#pragma unroll
            for (int b = 0; b < FB; b++) T[b * c + j] = acc[b];
        }
    }
}
// Qn = Z T : lanes = (row ii = lane & 7, column-part = lane >> 3); eight rows of Z per warp
// executed by warps w0 .. NW-1 only (w0 = 1 while warp 0 solves the Rayleigh-Ritz eigenproblem)
template <int NW>
__device__ __forceinline__ void mul_z_t(const double *Z, int r, int c, const double *T, double *Qn, int ldq, int w0 = 0) {
    const int warp = (threadIdx.x >> 5) - w0, lane = threadIdx.x & 31, ii = lane & 7, part = lane >> 3;
    if (warp < 0) return;
    for (int i0 = warp * 8; i0 < r; i0 += (NW - w0) * 8) {
        const int i = i0 + ii;
        const bool live = i < r;
        const double *zr = Z + (live ? i : 0);
        double acc[FB];
#pragma unroll
        for (int b = 0; b < FB; b++) acc[b] = 0.0;
        for (int j = part; j < c; j += 4) {
            const double z = live ? zr[(size_t)j * r] : 0.0;
#pragma unroll
            for (int b = 0; b < FB; b++) acc[b] = fma(z, T[b * c + j], acc[b]);
        }
#pragma unroll
        for (int b = 0; b < FB; b++) {
            acc[b] += __shfl_xor_sync(0xffffffffu, acc[b], 8);
            acc[b] += __shfl_xor_sync(0xffffffffu, acc[b], 16);
        }
        if (part == 0 && live) {
#pragma unroll
            for (int b = 0; b < FB; b++) Qn[b * ldq + i] = acc[b];
        }
    }
}

// Q <- Qn L^-T with Qn'Qn = L L' (Cholesky-QR); a column whose pivot collapses (numerically dependent) becomes zero.
// G: 36 + 8 doubles of shared scratch.  In-place (Qdst == Qsrc) is allowed.  Ends with a __syncthreads.
template <int NW>
__device__ void cholqr(const double *Qsrc, double *Qdst, int r, int ldq, double *G) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int e = warp; e < FB * (FB + 1) / 2; e += NW) {       // lower triangle, e = i(i+1)/2 + j
        int i = 0;
        while ((i + 1) * (i + 2) / 2 <= e) i++;
        const int j = e - i * (i + 1) / 2;
        const double *ci = Qsrc + i * ldq, *cj = Qsrc + j * ldq;
        double s = 0.0;
        for (int t = lane; t < r; t += 32) s = fma(ci[t], cj[t], s);
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) G[e] = s;
    }
    __syncthreads();
    if (warp == 0) {                                            // lane i owns row i of the factor
        const int i = lane & 7;
        double g[FB];
#pragma unroll
        for (int j = 0; j < FB; j++) g[j] = (j <= i) ? G[i * (i + 1) / 2 + j] : 0.0;
        const double d0 = g[i >= 0 ? 0 : 0];
        (void)d0;
        double inv_mine = 0.0;
#pragma unroll
        for (int j = 0; j < FB; j++) {
            const double orig = __shfl_sync(0xffffffffu, G[j * (j + 1) / 2 + j], 0);   // same value on every lane
            const double d = __shfl_sync(0xffffffffu, g[j], j);
            const bool dead = !(d > orig * 1e-22) || !(orig > 0.0);
            const double inv = dead ? 0.0 : rsqrt(d);
            const double lij = g[j] * inv;                       // L[i][j] for i >= j (sqrt(d) on the diagonal)
#pragma unroll
            for (int k = j + 1; k < FB; k++) {
                const double lkj = __shfl_sync(0xffffffffu, lij, k);
                g[k] = fma(-lij, lkj, g[k]);                     // meaningful for i >= k
            }
            g[j] = lij;
            if (i == j) inv_mine = inv;
        }
        __syncwarp();
        if (lane < FB) {
#pragma unroll
            for (int j = 0; j < FB; j++)
                if (j <= i) G[i * (i + 1) / 2 + j] = g[j];
            G[36 + i] = inv_mine;
        }
    }
    __syncthreads();
    if (tid < r) {
        double y[FB];
#pragma unroll
        for (int j = 0; j < FB; j++) {
            double s = Qsrc[j * ldq + tid];
#pragma unroll
            for (int a = 0; a < j; a++) s = fma(-G[j * (j + 1) / 2 + a], y[a], s);
            y[j] = s * G[36 + j];
        }
#pragma unroll
        for (int j = 0; j < FB; j++) Qdst[j * ldq + tid] = y[j];
    }
    __syncthreads();
}

// Symmetric 8 x 8 eigenproblem  G = V diag(theta) V'  by cyclic two-sided Jacobi inside ONE warp (Rayleigh-Ritz of the
// block).  Lane l owns row (l & 7) of the working matrix and of V (the four groups of eight lanes run the same
// computation).  A sweep is the 7-step round-robin tournament on 8 indices, 4 disjoint pairs per step; lane group m
// computes the rotation of pair m and broadcasts it; all register indices are compile-time constants.
// Pairs whose diagonal entries are BOTH below `cut` are left alone: pca_image keeps component k only while the
// cumulative variance is below pratio of the total, so a kept component carries at least (1 - pratio) total / min(r,c);
// with cut an eighth of that, the principal block of the "small" indices has trace < 7 cut and cannot hide a component
// that would be kept, and the sum of the Ritz values (all that is used of the small ones) is rotation-invariant.  Gt: lower triangle, e = i(i+1)/2 + j;  Vout[a * 8 + b] = V_ab.
static __device__ void eig8_warp(const double *Gt, double *Vout, double cut) {
    const int lane = threadIdx.x & 31, i = lane & 7;
    double a[FB], v[FB], dg;
#pragma unroll
    for (int j = 0; j < FB; j++) {
        a[j] = (j <= i) ? Gt[i * (i + 1) / 2 + j] : Gt[j * (j + 1) / 2 + i];
        v[j] = (i == j) ? 1.0 : 0.0;
    }
    dg = Gt[i * (i + 1) / 2 + i];
    const int grp = lane >> 3;                               // group m of eight lanes computes the rotation of pair m
    for (int sweep = 0; sweep < 16; sweep++) {
        int rotated = 0;
#pragma unroll
        for (int s = 0; s < 7; s++) {
            // pair m of step s: (7, s) for m = 0, ((s + m) % 7, (s - m) % 7) otherwise, ordered P < Q
            double apq = 0.0, app = 0.0, aqq = 0.0;
#pragma unroll
            for (int m = 0; m < 4; m++) {
                const int x = (m == 0) ? 7 : (s + m) % 7, y = (m == 0) ? s : (s - m + 7) % 7;
                const int P = x < y ? x : y, Q = x < y ? y : x;
                const double e = __shfl_sync(0xffffffffu, a[Q], P, 8);
                const double dp = __shfl_sync(0xffffffffu, dg, P, 8);
                const double dq = __shfl_sync(0xffffffffu, dg, Q, 8);
                if (grp == m) { apq = e; app = dp; aqq = dq; }
            }
            // t = tan(theta) from  t = h sgn(d) / (|d| + sqrt(d^2 + h^2)),  d = aqq - app, h = 2 apq.  Only c^2 + s^2 = 1
            // needs full accuracy (c = rsqrt(1 + t^2), s = t c), so the quotient uses rsqrt instead of IEEE div/sqrt.
            double c = 1.0, sg = 0.0, t = 0.0;
            if (apq * apq > (DBL_EPSILON * DBL_EPSILON) * fabs(app * aqq) && apq != 0.0 && (app >= cut || aqq >= cut)) {
                const double d = aqq - app, h = 2.0 * apq;
                const double w = fma(d, d, h * h);
                const double den = fabs(d) + w * rsqrt(w);
                const double ri = rsqrt(den);
                t = copysign(h, h * d) * (ri * ri);
                if (d == 0.0) t = copysign(1.0, apq);
                if (isfinite(t)) { c = rsqrt(fma(t, t, 1.0)); sg = c * t; } else t = 0.0;
            }
            double myc = 1.0, mys = 0.0, dnew = dg;
            bool amP = false;
#pragma unroll
            for (int m = 0; m < 4; m++) {
                const int x = (m == 0) ? 7 : (s + m) % 7, y = (m == 0) ? s : (s - m + 7) % 7;
                const int P = x < y ? x : y, Q = x < y ? y : x;
                const double cm = __shfl_sync(0xffffffffu, c, 8 * m);
                const double sm = __shfl_sync(0xffffffffu, sg, 8 * m);
                const double tp = __shfl_sync(0xffffffffu, t * apq, 8 * m);
                const double pp = __shfl_sync(0xffffffffu, app, 8 * m);
                const double qq = __shfl_sync(0xffffffffu, aqq, 8 * m);
                if (sm != 0.0) rotated = 1;
                if (i == P) { amP = true; myc = cm; mys = sm; dnew = pp - tp; }
                if (i == Q) { myc = cm; mys = sm; dnew = qq + tp; }
                // columns P, Q of this lane's rows (A J and V J)
                const double ap = a[P], aq = a[Q], vp = v[P], vq = v[Q];
                a[P] = cm * ap - sm * aq; a[Q] = sm * ap + cm * aq;
                v[P] = cm * vp - sm * vq; v[Q] = sm * vp + cm * vq;
            }
            // rows: J' (A J); the partner of index i in step s of the tournament
            const int partner = (i == 7) ? s : ((i == s) ? 7 : (2 * s - i + 14) % 7);
#pragma unroll
            for (int j = 0; j < FB; j++) {
                const double th = __shfl_sync(0xffffffffu, a[j], partner, 8);
                a[j] = amP ? (myc * a[j] - mys * th) : (mys * th + myc * a[j]);
            }
            dg = dnew;
        }
        if (!rotated) break;
    }
    if (lane < FB) {
#pragma unroll
        for (int j = 0; j < FB; j++) Vout[i * FB + j] = v[j];
    }
}

// Rayleigh-Ritz on the block: G = T'T (= Q' Z Z' Q), V from eig8_warp, then T <- T V, Q <- Q V and Qn <- (Z T) V.
// FGm: 36 doubles Gram + 64 doubles V.  Ends with a __syncthreads.
template <int NW>
__device__ void rayleigh_ritz(const double *Z, double *T, int c, double *Q, double *Qn, int r, int ldq, double *FGm, double cut) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    double *Gt = FGm, *V = FGm + 36;
    for (int e = warp; e < FB * (FB + 1) / 2; e += NW) {
        int i = 0;
        while ((i + 1) * (i + 2) / 2 <= e) i++;
        const int j = e - i * (i + 1) / 2;
        const double *ci = T + i * c, *cj = T + j * c;
        double s = 0.0;
        for (int t = lane; t < c; t += 32) s = fma(ci[t], cj[t], s);
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) Gt[e] = s;
    }
    __syncthreads();
    // warp 0 solves the 8 x 8 eigenproblem; the other warps meanwhile form Qn = Z T = Z Z' Q with the UNrotated block
    // (rotating afterwards commutes: Z Z' (Q V) = (Z Z' Q) V), which the residual test of the caller needs
    if (warp == 0) eig8_warp(Gt, V, cut);
    else mul_z_t<NW>(Z, r, c, T, Qn, ldq, 1);
    __syncthreads();
    for (int u = tid; u < 2 * r + c; u += nthr) {
        double *base = (u < r) ? Q + u : (u < r + c ? T + (u - r) : Qn + (u - r - c));
        const int ld = (u < r || u >= r + c) ? ldq : c;
        double x[FB], y[FB];
#pragma unroll
        for (int b = 0; b < FB; b++) { x[b] = base[b * ld]; y[b] = 0.0; }
#pragma unroll
        for (int aa = 0; aa < FB; aa++) {
#pragma unroll
            for (int b = 0; b < FB; b++) y[b] = fma(x[aa], V[aa * FB + b], y[b]);
        }
#pragma unroll
        for (int b = 0; b < FB; b++) base[b * ld] = y[b];
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// Fastest path: ONE component.  A star stamp over a noise floor has one principal component carrying > pratio of the
// variance (every star of the benchmark catalogues); then pca_image keeps exactly k = 1 and all that is needed is the
// leading left singular vector u and t = Z'u.  Plain power iteration  q <- Z Z' q / |.|  from the central column
// converges by (sigma_2 / sigma_1)^2 ~ 1e-5 per step and needs no small-matrix work at all (the block path spends 180k
// cycles per star in its Cholesky-QRs and 8 x 8 eigenproblem; this one ~10k).  It is ACCEPTED only when the answer is
// provably the full SVD's:  theta = |Z'q|^2 (a lower bound of sigma_1^2) already reaches pratio of |Z|_F^2 with the same
// 1e-11 guard band as choose_pcadim's comparison in the block path, and  |Z Z'q - theta q| <= 1e-13 |Z|_F^2.
// Anything else returns 0 with Z untouched and the block path takes over.  q, qn: r doubles; t: c doubles;
// sc2: 4 NW doubles of scratch.  On acceptance q = u (unit), t = Z'u, *theta_out = theta.
// ---------------------------------------------------------------------------------------------
template <int NW>
__device__ int power_pca(const double *Z, int r, int c, double fro, double pratio, double *q, double *qn, double *t,
                         double *sc2, double *theta_out, int *dbg_it) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    const int sub = lane & 7, part = lane >> 3;
    int buf = 0;
    double v2[2];
    {   // start: the central column, normalised
        const double *z0 = Z + (size_t)(c / 2) * r;
        double s = 0.0;
        for (int i = tid; i < r; i += nthr) { const double z = z0[i]; s = fma(z, z, s); }
        v2[0] = s; v2[1] = 0.0;
        cta_sum<2, NW>(v2, sc2, buf);
        if (!(v2[0] > 0.0) || !isfinite(v2[0])) return 0;
        const double inv = 1.0 / sqrt(v2[0]);
        for (int i = tid; i < r; i += nthr) q[i] = z0[i] * inv;
        __syncthreads();
    }
    const double den = (double)(c - 1), thres = (fro / den) * pratio, tol = 1e-13 * fro;
    for (int it = 1; it <= 8; it++) {
        // t = Z' q : eight columns of Z per warp, lanes = (column sub, row part)
        for (int j0 = warp * 8; j0 < c; j0 += NW * 8) {
            const int j = j0 + sub;
            const bool live = j < c;
            const double *zc = Z + (size_t)(live ? j : 0) * r;
            double a0 = 0.0, a1 = 0.0;
            int i = part;
            for (; i + 4 < r; i += 8) { a0 = fma(zc[i], q[i], a0); a1 = fma(zc[i + 4], q[i + 4], a1); }
            if (i < r) a0 = fma(zc[i], q[i], a0);
            double acc = live ? a0 + a1 : 0.0;
            acc += __shfl_xor_sync(0xffffffffu, acc, 8);
            acc += __shfl_xor_sync(0xffffffffu, acc, 16);
            if (part == 0 && live) t[j] = acc;
        }
        __syncthreads();
        // qn = Z t : eight rows of Z per warp, lanes = (row sub, column part); theta = |t|^2
        for (int i0 = warp * 8; i0 < r; i0 += NW * 8) {
            const int i = i0 + sub;
            const bool live = i < r;
            const double *zr = Z + (live ? i : 0);
            double a0 = 0.0, a1 = 0.0;
            int j = part;
            for (; j + 4 < c; j += 8) { a0 = fma(zr[(size_t)j * r], t[j], a0); a1 = fma(zr[(size_t)(j + 4) * r], t[j + 4], a1); }
            if (j < c) a0 = fma(zr[(size_t)j * r], t[j], a0);
            double acc = live ? a0 + a1 : 0.0;
            acc += __shfl_xor_sync(0xffffffffu, acc, 8);
            acc += __shfl_xor_sync(0xffffffffu, acc, 16);
            if (part == 0 && live) qn[i] = acc;
        }
        double th = 0.0;
        for (int j = tid; j < c; j += nthr) th = fma(t[j], t[j], th);
        v2[0] = th; v2[1] = 0.0;
        cta_sum<2, NW>(v2, sc2, buf);                               // also publishes qn
        const double theta = v2[0];
        double rs = 0.0, nn = 0.0;
        for (int i = tid; i < r; i += nthr) {
            const double x = qn[i], d = x - theta * q[i];
            rs = fma(d, d, rs);
            nn = fma(x, x, nn);
        }
        v2[0] = rs; v2[1] = nn;
        cta_sum<2, NW>(v2, sc2, buf);
        const double acc1 = theta / den;
        if (!(acc1 < thres) && !(fabs(acc1 - thres) <= 1e-11 * thres) && v2[0] <= tol * tol && theta > 0.0 && isfinite(fro)) {
            if (dbg_it) *dbg_it = 50 + it;
            *theta_out = theta;
            return 1;
        }
        // a Rayleigh quotient well below the threshold after two steps will not reach it: this is not a one-component star
        if (!(v2[1] > 0.0) || !isfinite(v2[1]) || (it >= 2 && acc1 < 0.97 * thres)) return 0;
        const double inv = 1.0 / sqrt(v2[1]);
        for (int i = tid; i < r; i += nthr) q[i] = qn[i] * inv;
        __syncthreads();
    }
    return 0;
}

// returns k >= 1 on acceptance (ford[0..k) = block columns of the kept components, Q = Ritz vectors, T = Z'Q), else 0
template <int NW, int RT>
__device__ int fast_pca(const double *Z, int r, int c, double fro, int ncomp, double pratio, double *Q, double *Qn,
                        double *T, int ldq, double *G, double *fth, int *ford, int *s_flag, int *dbg_it, int usecut, int first_check) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    const int j0 = c / 2 - FB / 2;
    for (int e = tid; e < FB * r; e += nthr) {
        const int b = e / r, i = e - b * r;
        Qn[b * ldq + i] = Z[i + (size_t)(j0 + b) * r];
    }
    __syncthreads();
    cholqr<NW>(Qn, Q, r, ldq, G);
    cholqr<NW>(Q, Q, r, ldq, G);
    // the small-pair cut of eig8_warp measured SLOWER on the bench catalogue (profiles/README.md); off unless debug & 4
    const double rr_cut = usecut ? (1.0 - pratio) * fro / (8.0 * (double)(r < c ? r : c)) : 0.0;
    for (int it = 1; it <= 12; it++) {
        mul_zt_q<NW>(Z, r, c, Q, ldq, T);
        __syncthreads();
        const bool check = (it == first_check || it == 3 || it == 5 || it == 8 || it == 12);
        if (check) {
            // Rayleigh-Ritz: T'T = V Theta V' (8 x 8, one warp), then T <- T V and Q <- Q V
            rayleigh_ritz<NW>(Z, T, c, Q, Qn, r, ldq, G + 64, rr_cut);
            for (int b = warp; b < FB; b += NW) {
                double s = 0.0;
                for (int j = lane; j < c; j += 32) s = fma(T[b * c + j], T[b * c + j], s);
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (lane == 0) fth[b] = s;
            }
            __syncthreads();
            if (tid == 0) {
                for (int b = 0; b < FB; b++) {                  // stable descending rank
                    int rank = 0;
                    for (int i = 0; i < FB; i++) rank += (fth[i] > fth[b]) || (fth[i] == fth[b] && i < b);
                    ford[rank] = b;
                }
                double sum8 = 0.0;
                for (int b = 0; b < FB; b++) sum8 += fth[ford[b]];
                double E = fro - sum8;
                if (E < 0.0) E = 0.0;
                const int m = r < c ? r : c;
                const double den = (double)(c - 1), thres = (fro / den) * pratio;
                int k = 1, ok = 1;
                double acc = fth[ford[0]] / den;
                for (;;) {                                       // choose_pcadim on the Ritz values
                    if (!(k < m && k < ncomp)) break;
                    if (fabs(acc - thres) <= 1e-11 * thres) { ok = 0; break; }
                    if (!(acc < thres)) break;
                    if (k >= FB_KMAX) { ok = 0; break; }
                    acc += fth[ford[k]] / den;
                    k++;
                }
                const double thk = fth[ford[k - 1]];
                if (!(thk > 0.0) || !(E <= 0.25 * thk) || !isfinite(fro) || !(thk >= 8.0 * rr_cut)) ok = 0;
                *s_flag = ok ? k : 0;
            }
            __syncthreads();
            const int k = *s_flag;
            if (k == 0) return 0;
            for (int a = warp; a < k; a += NW) {                 // residual |Z Z'u - theta u|^2 (Qn = Z Z' Q from rayleigh_ritz)
                const int b = ford[a];
                const double th = fth[b];
                double s = 0.0;
                for (int i = lane; i < r; i += 32) {
                    const double d = Qn[b * ldq + i] - th * Q[b * ldq + i];
                    s = fma(d, d, s);
                }
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (lane == 0) G[48 + a] = s;
            }
            __syncthreads();
            const double tol = 1e-13 * fro;
            bool conv = true;
            for (int a = 0; a < k; a++) conv = conv && (G[48 + a] <= tol * tol);
            if (conv) { if (dbg_it) *dbg_it = it; return k; }
            __syncthreads();
        } else {
            mul_z_t<NW>(Z, r, c, T, Qn, ldq);
            __syncthreads();
        }
        cholqr<NW>(Qn, Q, r, ldq, G);
        cholqr<NW>(Q, Q, r, ldq, G);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// smooth(data, a) (dataPreprocessing.jl:112-139) for a NaN-free image R (r x c, shared or global) -> out (global).
// The reference's kernel is the outer product kz (x) kz / S, so the renormalised correlation factorises:
//     out(i,j) = sum_dj kz(dj) [ sum_di kz(di) R(i+di, j+dj) ] / (A_i B_j),   A_i, B_j = sums of the taps inside the image.
// One thread owns row i of a column segment [j0, j1) and slides along j: the 2 LA + 1 vertical sums it needs live in a
// register window that rotates by full unrolling (no moves, no local memory), so a pixel costs 2 LA + 1 shared loads
// and 2 (2 LA + 1) FMAs instead of (2 LA + 1)^2 loads.  Lanes run along i: conflict-free shared reads, coalesced stores.
// ---------------------------------------------------------------------------------------------
// SH: R is known to lie in the CTA's dynamic shared memory; the pointer is re-derived from the shared base so that the
// (not inlined) body reads it with LDS instead of generic loads.
template <int LA, bool SH = false>
__device__ __noinline__ void smooth_sep(const double *Rg, double *out, int r, int c, const double *kz, int nthr) {
    constexpr int KW = 2 * LA + 1;
    const double *R = Rg;
    if constexpr (SH) {
        extern __shared__ __align__(16) unsigned char smem_sep[];
        R = reinterpret_cast<const double *>(smem_sep + (__cvta_generic_to_shared(Rg) - __cvta_generic_to_shared(smem_sep)));
    }
    int nseg = nthr / r;
    if (nseg < 1) nseg = 1;
    if (nseg > c) nseg = c;
    const int seglen = (c + nseg - 1) / nseg;
    double wz[KW], s1 = 0.0;
#pragma unroll
    for (int t = 0; t < KW; t++) { wz[t] = kz[t]; s1 += wz[t]; }
    const double inv_s1 = 1.0 / s1;
    for (int unit = threadIdx.x; unit < r * nseg; unit += nthr) {
        const int seg = unit / r, i = unit - seg * r;
        const int j0 = seg * seglen, j1 = (j0 + seglen < c) ? j0 + seglen : c;
        double wv[KW], Ai = 0.0;
        int ro[KW];                                   // clamped row offsets (out-of-image rows carry weight 0)
#pragma unroll
        for (int t = 0; t < KW; t++) {
            const int ii = i + t - LA;
            const bool ok = ii >= 0 && ii < r;
            wv[t] = ok ? wz[t] : 0.0;
            ro[t] = ok ? ii : i;
            Ai += wv[t];
        }
        const double inv_ai = 1.0 / Ai;
        double h[KW];
#pragma unroll
        for (int t = 0; t < KW - 1; t++) {            // prime the window with columns j0 - LA .. j0 + LA - 1
            const int jj = j0 - LA + t;
            double vs = 0.0;
            if (jj >= 0 && jj < c) {
                const double *col = R + (size_t)jj * r;
#pragma unroll
                for (int d = 0; d < KW; d++) vs = fma(wv[d], col[ro[d]], vs);
            }
            h[t] = vs;
        }
        h[KW - 1] = 0.0;
        for (int jb = j0; jb < j1; jb += KW) {
#pragma unroll
            for (int u = 0; u < KW; u++) {
                const int j = jb + u;
                if (j < j1) {
                    const int jn = j + LA;
                    double vs = 0.0;
                    if (jn < c) {
                        const double *col = R + (size_t)jn * r;
#pragma unroll
                        for (int d = 0; d < KW; d++) vs = fma(wv[d], col[ro[d]], vs);
                    }
                    h[(u + KW - 1) % KW] = vs;
                    double acc = 0.0;
#pragma unroll
                    for (int t = 0; t < KW; t++) acc = fma(wz[t], h[(u + t) % KW], acc);
                    double inv_b = inv_s1;
                    if (j < LA || j >= c - LA) {
                        double Bj = 0.0;
#pragma unroll
                        for (int t = 0; t < KW; t++) { const int jj = j + t - LA; if (jj >= 0 && jj < c) Bj += wz[t]; }
                        inv_b = 1.0 / Bj;
                    }
                    out[i + (size_t)j * r] = acc * inv_ai * inv_b;
                }
            }
        }
    }
}
// rows handled by one lane in the Jacobi kernels: i = lane + 32 t, t < RT  (32 RT >= max(r, c))
// GW: the W / Y buffers live in the per-CTA global scratch (stamps too large for shared memory).  With GW = false every
// pointer derived from W is provably a shared-memory address, so the compiler emits LDS/STS with 32-bit addressing
// instead of generic LD/ST with 64-bit address arithmetic.
template <int NW, int RT, bool GW>
__global__ void __launch_bounds__(NW * 32, 1) pca_kernel(const PcaArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int r = a.r, c = a.c;
    const long long rc = (long long)r * c;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    const int kmaxp = (a.kmax + 1) & ~1, ldq = (r + 1) & ~1;
    // shared layout: [ misc 192 B | scratch 2*NW | mu r | sv c | fth 8 | ord c' ints | uoff, yoff kmax' ints | ford 8 ints |
    //                  (W rc | region) unless global ];  det (3 r doubles) sits between fth and ord;  region = Y (kmax x c)  overlaid with  Q, Qn (8 x ldq), T (8 x c), G 64
    long long *s_next = reinterpret_cast<long long *>(smem_raw);
    int *s_flag = reinterpret_cast<int *>(smem_raw + 16);
    unsigned long long *prof = (a.debug & 16) ? reinterpret_cast<unsigned long long *>(smem_raw + 64) : nullptr;   // 16 ticks
    double *scratch = reinterpret_cast<double *>(smem_raw + 192);
    double *mu = scratch + 2 * NW;
    double *sv = mu + r;
    double *fth = sv + c;
    double *det = fth + FB;                                       // d, e, tau of the direct path (3 r)
    int *ord = reinterpret_cast<int *>(det + 3 * r);
    int *uoff = ord + ((c + 1) & ~1);
    int *yoff = uoff + kmaxp;
    int *ford = yoff + kmaxp;
    double *tail = reinterpret_cast<double *>(ford + FB);
    double *W;
    if constexpr (GW) W = a.wglobal + (size_t)blockIdx.x * a.wstride;
    else W = tail;
    double *Y = W + rc;
    double *FQ = Y, *FQn = FQ + FB * ldq, *FT = FQn + FB * ldq, *FG = FT + FB * c;
    int buf = 0;

    const long long ntot = a.list ? (long long)*a.nlist : a.N;
    if (tid == 0) *s_next = (long long)atomicAdd(a.counter, 1ULL);
    for (;;) {
        __syncthreads();
        const long long work = *s_next;                          // position in the work list
        if (work >= ntot) break;
        const long long star = a.list ? a.list[work] : work;
        __syncthreads();                                         // everyone has read s_next before it is replaced
        // the next star's index is fetched now and published at the end of this one: the atomic's latency is hidden
        long long nxt = 0;
        if (tid == 0) nxt = (long long)atomicAdd(a.counter, 1ULL);
        const double *X = a.stamps + star * rc;
        double *out = a.pgf + star * rc;
        if (prof && tid == 0) { for (int i = 0; i < 16; i++) prof[i] = 0; prof[0] = clock64(); }

        // ---- load, nanToZero, finiteness (LAPACK raises on Inf -> shopt.jl:305-309 zeros + failedStars) ----
        int bad = 0, nrep = 0;
        for (long long p = tid; p < rc; p += nthr) {
            const double x0 = __ldg(X + p);
            const double x = nan_to_zero(x0);
            nrep += isnan(x0);
            if (!isfinite(x)) bad = 1;
            W[p] = x;
        }
        bad = __syncthreads_or(bad);
        if (bad) {
            for (long long p = tid; p < rc; p += nthr) out[p] = 0.0;
            if (tid == 0) { if (a.k_out) a.k_out[star] = 0; if (a.status) a.status[star] = SHOPT_PCA_FAILED; *s_next = nxt; }
            continue;
        }
        // ---- mean over the columns, centre.  Thread = (row i, column segment): lanes run along i (conflict-free shared
        //      reads), the segment sums of a row are added in segment order (the reference's mean(X, dims=2) adds the
        //      columns in order; the two differ in the last bit of the mean only). ----
        int nsg0 = nthr / r;
        if (nsg0 < 1) nsg0 = 1;
        if (nsg0 > c) nsg0 = c;
        if (nsg0 > 2 * FB) nsg0 = 2 * FB;                       // the segment sums sit in the block path's Q buffers
        const int sgl0 = (c + nsg0 - 1) / nsg0;
        for (int unit = tid; unit < r * nsg0; unit += nthr) {
            const int seg = unit / r, i = unit - seg * r;
            const int j0 = seg * sgl0, j1 = (j0 + sgl0 < c) ? j0 + sgl0 : c;
            double s = 0.0;
            const double *w = W + i + (size_t)j0 * r;
            for (int j = j0; j < j1; j++, w += r) s += *w;
            Y[seg * r + i] = s;
        }
        __syncthreads();
        for (int i = tid; i < r; i += nthr) {
            double s = Y[i];
            for (int seg = 1; seg < nsg0; seg++) s += Y[seg * r + i];
            mu[i] = s / c;
        }
        __syncthreads();
        double fro = 0.0;
        for (int unit = tid; unit < r * nsg0; unit += nthr) {
            const int seg = unit / r, i = unit - seg * r;
            const int j0 = seg * sgl0, j1 = (j0 + sgl0 < c) ? j0 + sgl0 : c;
            const double mui = mu[i];
            double *w = W + i + (size_t)j0 * r;
            for (int j = j0; j < j1; j++, w += r) {
                const double z = *w - mui;
                *w = z;
                fro = fma(z, z, fro);
            }
        }
        fro = block_sum1<NW>(fro, scratch, buf);                // |Z|_F^2; the block sum also publishes W
        // Masked pixels become 0.001 (masks.jl:43), each an outlier in a direction of its own.  When their energy alone
        // exceeds half of the variance the kept components may leave out, k runs into the tens and the subspace
        // iterations below would only be abandoned (205k cycles per star on the masked BASELINE catalogue): such stars
        // go straight to the direct path.  A scheduling decision only: both paths are checked a posteriori.
        const bool direct_ok = !GW && a.eigscratch && r <= c && r >= 16 && r <= nthr && r <= 32 * RT;
        const double nrep_all = (a.debug & 64) ? 0.0 : block_sum1<NW>((double)nrep, scratch, buf);
        const bool flat = direct_ok && nrep_all * 1e-6 > 0.5 * (1.0 - a.pratio) * fro;

        if (prof && tid == 0) prof[1] = clock64();
        // ---- fast path: block subspace iteration with a-posteriori acceptance ----
        int k = 0, status = SHOPT_PCA_OK, dbg_it = 0;
        if (!flat && a.fast && r >= 2 * FB && c >= 2 * FB && r <= nthr && a.kmax >= 1 && !(a.debug & 32)) {
            // one component: q, qn, t in the block path's buffers (column 0 of Q, Qn, T); scratch in its Gram area
            k = power_pca<NW>(W, r, c, fro, a.pratio, FQ, FQn, FT, FG, fth, &dbg_it);
            if (k == 1 && tid == 0) ford[0] = 0;
            __syncthreads();
        }
        if (k == 0 && !flat && a.fast && r >= 2 * FB && c >= 2 * FB && r <= nthr)
            k = fast_pca<NW, RT>(W, r, c, fro, a.ncomp, a.pratio, FQ, FQn, FT, ldq, FG, fth, ford, s_flag, &dbg_it, a.debug & 4, (a.debug & 8) ? 2 : 3);
        const bool fastk = k > 0;
        if (prof && tid == 0) prof[2] = clock64();
        bool eigk = false;
        if (!fastk && direct_ok) {
            // ---- direct path: eigenvectors of the r x r Gram matrix (pca_eig.cuh); 0 = a check failed ----
            double *Cg = a.eigscratch + (size_t)blockIdx.x * a.eigstride;
            k = gram_eig_pca<NW, RT>(W, r, c, a.ncomp, a.pratio, Y, det, sv, fth, s_flag, Cg, Cg + (size_t)r * r, prof);
            eigk = k > 0;
            dbg_it = eigk ? 100 : 200;
            if (!eigk) {                                         // W was consumed: centre the stamp again
                for (long long p = tid; p < rc; p += nthr) W[p] = nan_to_zero(__ldg(X + p)) - mu[(int)(p % r)];
                __syncthreads();
            }
        }
        if (fastk) {
            if (tid < k) {
                uoff[tid] = (int)(FQ - W) + ford[tid] * ldq;
                yoff[tid] = (int)(FT - W) + ford[tid] * c;
            }
            __syncthreads();
        } else if (eigk) {
            if (tid < k) { uoff[tid] = (int)rc + tid * r; yoff[tid] = tid * c; }     // U_k in the region, Y into W
            __syncthreads();
        } else {
            // ---- full SVD: one-sided Jacobi on all c columns; the columns become sigma_j u_j ----
            // columns whose squared norm falls below eps^2 |Z|_F^2 are numerically zero (always c - r + 1 of them when
            // c > r): rotating them only shuffles round-off, so they are left alone and do not hold up convergence
            int sweeps;
            const int rotating = jacobi_columns<NW, RT>(W, r, c, r, nullptr, 0, 0, fro * (DBL_EPSILON * DBL_EPSILON), 60, sweeps);
            if (rotating && sweeps >= 60) status = SHOPT_PCA_NOCONV;
            // principal variances, stable descending rank, choose k (choose_pcadim)
            for (int j = tid; j < c; j += nthr) {
                double s = 0.0;
                const double *w = W + (size_t)j * r;
                for (int i = 0; i < r; i++) s = fma(w[i], w[i], s);
                sv[j] = s;
            }
            __syncthreads();
            for (int j = tid; j < c; j += nthr) {
                const double x = sv[j];
                int rank = 0;
                for (int i = 0; i < c; i++) rank += (sv[i] > x) || (sv[i] == x && i < j);
                ord[rank] = j;
            }
            __syncthreads();
            if (tid == 0) {
                const int m = r < c ? r : c;
                const double den = (double)(c - 1);
                double vsum = 0.0;
                for (int j = 0; j < m; j++) vsum += sv[ord[j]] / den;
                int kk = 1;
                double acc = sv[ord[0]] / den;
                const double thres = vsum * a.pratio;
                while (kk < m && kk < a.ncomp && acc < thres) { acc += sv[ord[kk]] / den; kk++; }
                *s_flag = kk;
            }
            __syncthreads();
            k = *s_flag;
            // normalise the k leading columns: u_a = w_a / sigma_a
            for (int e = tid; e < k * r; e += nthr) {
                const int aa = e / r, i = e - aa * r;
                const int col = ord[aa];
                const double nrm = sqrt(sv[col]);
                if (nrm > 0.0) W[i + (size_t)col * r] /= nrm;
            }
            if (tid < k) { uoff[tid] = ord[tid] * r; yoff[tid] = (int)rc + tid * c; }
            __syncthreads();
        }
        if (prof && tid == 0) prof[9] = clock64();
        if (!fastk) {
            // Y = U_k' Z, Z re-read from the stamp (transform); a warp holds its column of Z in registers
            for (int j = warp; j < c; j += NW) {
                double zt[RT];
#pragma unroll
                for (int t = 0; t < RT; t++) {
                    const int i = lane + 32 * t;
                    zt[t] = (i < r) ? nan_to_zero(__ldg(X + i + (size_t)j * r)) - mu[i] : 0.0;
                }
                for (int aa = 0; aa < k; aa++) {
                    const double *ua = W + uoff[aa];
                    double s = 0.0;
#pragma unroll
                    for (int t = 0; t < RT; t++) {
                        const int i = lane + 32 * t;
                        if (i < r) s = fma(ua[i], zt[t], s);
                    }
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                    if (lane == 0) W[yoff[aa] + j] = s;
                }
            }
            __syncthreads();
        }
        if (prof && tid == 0) prof[10] = clock64();
        // the next star's stamp starts its way into L2 now (one bulk prefetch); its load phase then hits L2
        if (tid == 0 && nxt < ntot && !(a.debug & 2)) {
            const long long nstar = a.list ? a.list[nxt] : nxt;
            const uintptr_t p0 = (uintptr_t)(a.stamps + nstar * rc), p1 = (uintptr_t)(a.stamps + (nstar + 1) * rc);
            const uintptr_t lo = p0 & ~(uintptr_t)15, hi = p1 & ~(uintptr_t)15;      // stays inside the batch
            if (hi > lo)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(lo), "r"((uint32_t)(hi - lo)) : "memory");
        }
        // ---- reconstruct: R = U_k Y + mean.  Thread = (row i, column segment); the fast path leaves U and Y outside W,
        //      so R goes straight into W when it is smoothed next; the Jacobi path (U inside W) goes through `out` ----
        const bool smoothing = !a.unity_sum && a.lanczos_a > 0;
        // ---- smooth(reconstruction) without ever forming the reconstruction.  R = mean 1' + U_k Y is a sum of k + 1 outer
        //      products and smooth() (dataPreprocessing.jl:112-139, NaN-free image) is a separable linear map with a
        //      separable normalisation:  out(i,j) = sum_di sum_dj kz(di) kz(dj) R(i+di, j+dj) / (A_i B_j),  A_i, B_j = sums of
        //      the taps inside the image.  Hence  out = M 1' + U' Y'  with the 1-D smoothed factors
        //          M(i) = sum_di kz(di) mean(i+di) / A_i,  U'_a(i) = sum_di kz(di) u_a(i+di) / A_i,  Y'_a(j) = sum_dj kz(dj) y_a(j+dj) / B_j
        //      -- the same algebra, k FMAs per pixel instead of a reconstruction pass plus a 2 (2a+1)-tap stencil pass
        //      (51 % of the instructions of a one-component star).  W is free by now on these two paths. ----
        bool fused = false;
        if (smoothing && (fastk || eigk) && a.lanczos_a <= 7 &&
            (size_t)(eigk ? k * c : 0) + (size_t)(k + 1) * r + (size_t)k * c <= (size_t)rc) {
            const int la = a.lanczos_a, kw = 2 * la + 1;
            double *Ms = W + (eigk ? k * c : 0), *Us = Ms + r, *Ys = Us + (size_t)k * r;
            int nf = 0;
            for (int e = tid; e < (k + 1) * r; e += nthr) {          // vertical factors: the mean and the k columns of U
                const int aa = e / r, i = e - aa * r;
                const double *src = (aa == 0) ? mu : W + uoff[aa - 1];
                double acc = 0.0, A = 0.0;
                for (int t = 0; t < kw; t++) {
                    const int ii = i + t - la;
                    if (ii >= 0 && ii < r) { acc = fma(a.kz[t], src[ii], acc); A += a.kz[t]; }
                }
                const double val = acc / A;
                if (!isfinite(val)) nf = 1;
                Ms[e] = val;
            }
            for (int e = tid; e < k * c; e += nthr) {                // horizontal factors: the k rows of Y
                const int aa = e / c, j = e - aa * c;
                const double *src = W + yoff[aa];
                double acc = 0.0, B = 0.0;
                for (int t = 0; t < kw; t++) {
                    const int jj = j + t - la;
                    if (jj >= 0 && jj < c) { acc = fma(a.kz[t], src[jj], acc); B += a.kz[t]; }
                }
                const double val = acc / B;
                if (!isfinite(val)) nf = 1;
                Ys[e] = val;
            }
            nf = __syncthreads_or(nf);                               // also publishes the factors
            if (!nf) {
                fused = true;
                int nsg = nthr / r;
                if (nsg < 1) nsg = 1;
                if (nsg > c) nsg = c;
                const int sgl = (c + nsg - 1) / nsg;
                for (int unit = tid; unit < r * nsg; unit += nthr) {
                    const int seg = unit / r, i = unit - seg * r;
                    const int j0 = seg * sgl, j1 = (j0 + sgl < c) ? j0 + sgl : c;
                    const double m = Ms[i];
                    double *o = out + i + (size_t)j0 * r;
                    if (k == 1) {
                        const double u0 = Us[i];
                        for (int j = j0; j < j1; j++, o += r) *o = fma(u0, Ys[j], m);
                    } else if (k == 2) {
                        const double u0 = Us[i], u1 = Us[r + i];
                        for (int j = j0; j < j1; j++, o += r) *o = fma(u1, Ys[c + j], fma(u0, Ys[j], m));
                    } else {
                        for (int j = j0; j < j1; j++, o += r) {
                            double sacc = m;
                            for (int aa = 0; aa < k; aa++) sacc = fma(Us[(size_t)aa * r + i], Ys[(size_t)aa * c + j], sacc);
                            *o = sacc;
                        }
                    }
                }
            }
        }
        if (!fused) {
        double *Rdst = (fastk && smoothing) ? W : out;
        int nseg = nthr / r;
        if (nseg < 1) nseg = 1;
        if (nseg > c) nseg = c;
        const int seglen = (c + nseg - 1) / nseg;
        double part = 0.0;
        int nonfinite = 0;
        for (int unit = tid; unit < r * nseg; unit += nthr) {
            const int seg = unit / r, i = unit - seg * r;
            const int j0 = seg * seglen, j1 = (j0 + seglen < c) ? j0 + seglen : c;
            const double mui = mu[i];
            if (k <= FB_KMAX) {
                double ua[FB_KMAX];
#pragma unroll
                for (int aa = 0; aa < FB_KMAX; aa++) ua[aa] = (aa < k) ? W[uoff[aa] + i] : 0.0;
                for (int j = j0; j < j1; j++) {
                    double sacc = 0.0;
#pragma unroll
                    for (int aa = 0; aa < FB_KMAX; aa++)
                        if (aa < k) sacc = fma(ua[aa], W[yoff[aa] + j], sacc);
                    sacc += mui;
                    if (!isfinite(sacc)) nonfinite = 1;
                    Rdst[i + (size_t)j * r] = sacc;
                    part += sacc;
                }
            } else {
                for (int j = j0; j < j1; j++) {
                    double sacc = 0.0;
                    for (int aa = 0; aa < k; aa++) sacc = fma(W[uoff[aa] + i], W[yoff[aa] + j], sacc);
                    sacc += mui;
                    if (!isfinite(sacc)) nonfinite = 1;
                    Rdst[i + (size_t)j * r] = sacc;
                    part += sacc;
                }
            }
        }
        if (a.unity_sum) {                                   // shopt.jl:295: pca ./ sum(pca), no smoothing
            const double tot = block_sum1<NW>(part, scratch, buf);
            for (int unit = tid; unit < r * nseg; unit += nthr) {        // same thread wrote these pixels
                const int seg = unit / r, i = unit - seg * r;
                const int j0 = seg * seglen, j1 = (j0 + seglen < c) ? j0 + seglen : c;
                for (int j = j0; j < j1; j++) out[i + (size_t)j * r] = out[i + (size_t)j * r] / tot;
            }
        } else if (smoothing) {                              // shopt.jl:299: smooth(pca, lanczos)
            nonfinite = __syncthreads_or(nonfinite);         // also publishes R
            if (!fastk) {
                for (int unit = tid; unit < r * nseg; unit += nthr) {    // same thread wrote these pixels
                    const int seg = unit / r, i = unit - seg * r;
                    const int j0 = seg * seglen, j1 = (j0 + seglen < c) ? j0 + seglen : c;
                    for (int j = j0; j < j1; j++) W[i + (size_t)j * r] = out[i + (size_t)j * r];
                }
                __syncthreads();
            }
            if (!nonfinite && a.lanczos_a == 3) {            // the shipped default (shopt.yml lanczos: 3)
                smooth_sep<3, !GW>(W, out, r, c, a.kz, nthr);
            } else {                                         // NaN-aware renormalisation, tap by tap like the reference
                const int la = a.lanczos_a, kw = 2 * la + 1;
                for (long long p = tid; p < rc; p += nthr) {
                    const int j = (int)(p / r), i = (int)(p - (long long)j * r);
                    if (isnan(W[p])) { out[p] = W[p]; continue; }
                    const int dlo = i - la < 0 ? 0 : i - la, dhi = i + la > r - 1 ? r - 1 : i + la;
                    const int elo = j - la < 0 ? 0 : j - la, ehi = j + la > c - 1 ? c - 1 : j + la;
                    double norm = 0.0, val = 0.0;
                    for (int di = dlo; di <= dhi; di++)
                        for (int dj = elo; dj <= ehi; dj++) {
                            const double d = W[di + (size_t)dj * r];
                            if (!isnan(d)) {
                                const double kk = a.kern[(di - i + la) + (dj - j + la) * kw];
                                val += kk * d;
                                norm += kk;
                            }
                        }
                    out[p] = val / norm;
                }
            }
        }
        }   // !fused
        if (prof) {
            __syncthreads();
            if (tid == 0) {
                const unsigned long long t_end = clock64();
                unsigned long long prev = prof[0];
                for (int i = 1; i <= 10; i++)
                    if (prof[i]) { atomicAdd(&g_pca_phase[i], prof[i] - prev); prev = prof[i]; }
                atomicAdd(&g_pca_phase[11], t_end - prev);
                atomicAdd(&g_pca_phase[0], 1ULL);
            }
        }
        if (tid == 0) {
            if (a.k_out) a.k_out[star] = k;
            if (a.status) a.status[star] = (a.debug & 1) ? (status | (dbg_it << 8)) : status;
            *s_next = nxt;
        }
    }
}

#ifndef SHOPT_PCA_RT   // the non-template kernels and their launchers live in the main translation unit only
// ---------------------------------------------------------------------------------------------
// smooth(data, a) alone (dataPreprocessing.jl:112-139); mode "smooth" applies it to nanToZero(star) (shopt.jl:316-348).
// One CTA per star; taps come straight from global memory (the 8 r c bytes of a stamp stay in L1/L2).
// ---------------------------------------------------------------------------------------------
struct SmoothArgs {
    const double *in;
    double *out;
    long long N;
    int r, c, la, nan_to_zero, use_smem;
    double kern[15 * 15];
    double kz[15];
};
// Persistent CTAs, grid-stride over stars.  When the stamp fits shared memory it is staged there (nanToZero applied on
// the way); a stamp without NaN/Inf then takes the separable sliding-window path, anything else the NaN-aware
// renormalised tap loop of the reference.  Larger stamps (301 x 301) read their taps from global memory (L1/L2).
__global__ void __launch_bounds__(512) smooth_kernel(const SmoothArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *skern = reinterpret_cast<double *>(smem_raw);          // 225 + 15 doubles
    double *skz = skern + 225;
    double *R = skz + 15;
    const int r = a.r, c = a.c, la = a.la, kw = 2 * la + 1;
    const long long rc = (long long)r * c;
    for (int t = threadIdx.x; t < 225; t += 512) skern[t] = (t < kw * kw) ? a.kern[t] : 0.0;
    if (threadIdx.x < 15) skz[threadIdx.x] = a.kz[threadIdx.x];
    for (long long star = blockIdx.x; star < a.N; star += gridDim.x) {
        const double *X = a.in + star * rc;
        double *out = a.out + star * rc;
        __syncthreads();
        int bad = 0;
        if (a.use_smem) {
            for (long long p = threadIdx.x; p < rc; p += 512) {
                double x = __ldg(X + p);
                if (a.nan_to_zero) x = nan_to_zero(x);
                if (!isfinite(x)) bad = 1;
                R[p] = x;
            }
            bad = __syncthreads_or(bad);
            if (la <= 0) {
                for (long long p = threadIdx.x; p < rc; p += 512) out[p] = R[p];
                continue;
            }
            if (!bad && la == 3) { smooth_sep<3, true>(R, out, r, c, skz, 512); continue; }
        }
        const double *src = a.use_smem ? R : X;
        const bool ntz = a.nan_to_zero && !a.use_smem;
        for (long long p = threadIdx.x; p < rc; p += 512) {
            const int j = (int)(p / r), i = (int)(p - (long long)j * r);
            double x = src[p];
            if (ntz) x = nan_to_zero(x);
            if (isnan(x) || la <= 0) { out[p] = x; continue; }
            const int dlo = i - la < 0 ? 0 : i - la, dhi = i + la > r - 1 ? r - 1 : i + la;
            const int elo = j - la < 0 ? 0 : j - la, ehi = j + la > c - 1 ? c - 1 : j + la;
            double norm = 0.0, val = 0.0;
            for (int di = dlo; di <= dhi; di++)
                for (int dj = elo; dj <= ehi; dj++) {
                    double d = src[di + (size_t)dj * r];
                    if (ntz) d = nan_to_zero(d);
                    if (!isnan(d)) {
                        const double kk = skern[(di - i + la) + (dj - j + la) * kw];
                        val += kk * d;
                        norm += kk;
                    }
                }
            out[p] = val / norm;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Pixel-grid fits that are element-wise per star (one CTA per star, grid-stride over stars):
//   mode "smooth" with unity_sum (shopt.jl:325-331): the 1-argument smooth() call raises, so the catch branch gives
//       nanToZero(star) ./ sum(nanToZero(star));
//   mode "chisq" (chisq.jl:1-11, shopt.jl:194-224): LBFGS over the r c pixel values of
//       sum(isfinite(t) ? t : 0),  t = (p - star)^2 / weight.  The objective is separable: a pixel whose term is a
//       finite positive-definite quadratic has the unique minimiser p = star; a pixel whose term is never finite or
//       has no curvature (weight 0, Inf or NaN, star not finite) has zero gradient and keeps its START value
//       (rand(r*c) in the reference; an explicit input here, `fill` when no start cube is given).
// ---------------------------------------------------------------------------------------------
struct ElemArgs {
    const double *in, *w, *x0;
    double *out;
    long long rc, N;
    double fill;
    int chisq, unity;
};
__global__ void __launch_bounds__(256) elem_pgf_kernel(const ElemArgs a) {
    __shared__ double sred[2 * 8];
    int buf = 0;
    for (long long star = blockIdx.x; star < a.N; star += gridDim.x) {
        const double *X = a.in + star * a.rc;
        double *out = a.out + star * a.rc;
        double part = 0.0;
        for (long long p = threadIdx.x; p < a.rc; p += 256) {
            double y;
            if (a.chisq) {
                const double d = __ldg(X + p), w = __ldg(a.w + star * a.rc + p);
                const bool live = isfinite(d) && isfinite(w) && w > 0.0;
                y = live ? d : (a.x0 ? __ldg(a.x0 + star * a.rc + p) : a.fill);
            } else {
                y = nan_to_zero(__ldg(X + p));
            }
            out[p] = y;
            part += y;
        }
        if (a.unity) {
            const double tot = block_sum1<8>(part, sred, buf);
            for (long long p = threadIdx.x; p < a.rc; p += 256) out[p] = out[p] / tot;   // same thread wrote it
        }
    }
}

int elem_pgf_launch(const double *d_in, const double *d_w, const double *d_x0, double fill, long long rc, long long N,
                    int chisq, int unity, double *d_out, cudaStream_t st) {
    if (!d_in || !d_out || rc < 1 || N < 0 || (chisq && !d_w)) return SHOPT_E_BADARG;
    if (N == 0) return SHOPT_OK;
    ElemArgs a;
    a.in = d_in; a.w = d_w; a.x0 = d_x0; a.out = d_out; a.rc = rc; a.N = N; a.fill = fill; a.chisq = chisq; a.unity = unity;
    long long grid = (long long)sm_count() * 8;
    if (grid > N) grid = N;
    elem_pgf_kernel<<<(unsigned)grid, 256, 0, st>>>(a);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

#endif  // !SHOPT_PCA_RT

// dataPreprocessing.jl:94-110: a*sinc(pi x)*sinc(pi x / a) at integer offsets, outer product, normalised by the sum
static void lanczos_kernel2d(int a, double *kernel, double *kz = nullptr) {
    const int w = 2 * a + 1;
    auto sinc = [](double x) { return x == 0 ? 1.0 : sin(x) / x; };
    auto lz = [&](double x) { return (fabs(x) >= a) ? 0.0 : (x == 0 ? 1.0 : a * sinc(M_PI * x) * sinc(M_PI * x / a)); };
    double sum = 0.0;
    for (int j = 0; j < w; j++)
        for (int i = 0; i < w; i++) kernel[i + j * w] = lz(i - a) * lz(j - a);
    for (int i = 0; i < w * w; i++) sum += kernel[i];
    for (int i = 0; i < w * w; i++) kernel[i] /= sum;
    if (kz) for (int i = 0; i < w; i++) kz[i] = lz(i - a);
}

// ---- launch -----------------------------------------------------------------------------------
static size_t pca_fixed_smem(int nw, int r, int c, int kmax) {
    return 192 + (size_t)2 * nw * 8 + (size_t)r * 8 + (size_t)c * 8 + FB * 8 + (size_t)3 * r * 8 + (size_t)((c + 1) & ~1) * 4 +
           (size_t)2 * ((kmax + 1) & ~1) * 4 + FB * 4;
}
static size_t pca_w_doubles(int r, int c, int kmax) {      // W plus the region shared by Y and the fast-path buffers
    const size_t y = (size_t)kmax * c, f = (size_t)2 * FB * ((r + 1) & ~1) + (size_t)FB * c + 192;
    return (size_t)r * c + (y > f ? y : f);
}
#ifndef SHOPT_PCA_RT
static size_t pca_heavy_scratch_bytes(int r, int c, int ncomp) {
    int kmax = ncomp < r ? ncomp : r; if (c < kmax) kmax = c; if (kmax < 1) kmax = 1;
    const size_t wdoubles = pca_w_doubles(r, c, kmax);
    if (pca_fixed_smem(32, r, c, kmax) + wdoubles * 8 <= 227 * 1024)
        return (r <= c && r >= 16) ? (size_t)sm_count() * 2 * r * r * 8 : 0;       // direct path: Gram matrix + reflectors
    return (size_t)sm_count() * wdoubles * 8;
}
// work list of the stars pca_power_kernel hands over (N indices) + its two counters, in front of pca_kernel's scratch
static size_t pca_list_bytes(long long N) { return N > 0 ? (((size_t)N * 8 + 64 + 255) & ~(size_t)255) : 0; }
size_t pca_scratch_bytes(int r, int c, int ncomp, long long N) { return pca_list_bytes(N) + pca_heavy_scratch_bytes(r, c, ncomp); }
#endif

template <int NW, int RT>
static int pca_launch_t(PcaArgs a, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    const size_t wdoubles = pca_w_doubles(a.r, a.c, a.kmax);
    size_t smem = pca_fixed_smem(NW, a.r, a.c, a.kmax) + wdoubles * 8;
    long long grid = sm_count();
    a.wglobal = nullptr; a.wstride = wdoubles;
    if (smem > 227 * 1024) {
        smem = pca_fixed_smem(NW, a.r, a.c, a.kmax);
        if ((size_t)grid * wdoubles * 8 > scratch_bytes) return SHOPT_E_NOMEM;
        a.wglobal = scratch;
    } else if (a.eigstride && scratch && (size_t)grid * a.eigstride * 8 <= scratch_bytes) {
        a.eigscratch = scratch;
    }
    void (*kern)(const PcaArgs) = nullptr;
    if (a.wglobal) {
        if constexpr (RT >= 5) kern = pca_kernel<NW, RT, true>;
    } else {
        if constexpr (RT <= 5) kern = pca_kernel<NW, RT, false>;
    }
    if (!kern) return SHOPT_E_UNSUPPORTED;                       // unreachable: 32 RT >= max(r, c) fixes which layout fits
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (grid > a.N) grid = a.N;
    SHOPT_CUDA_TRY(cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), st));
    if (a.debug & 16) {
        unsigned long long zero[16] = {0};
        SHOPT_CUDA_TRY(cudaMemcpyToSymbolAsync(g_pca_phase, zero, sizeof(zero), 0, cudaMemcpyHostToDevice, st));
    }
    kern<<<(unsigned)grid, NW * 32, smem, st>>>(a);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    if (a.debug & 16) {
        static const char *names[12] = {"stars", "load+centre", "fast path", "gram+copy", "tridiagonalise", "eigenvalues",
                                        "inverse iteration", "orthogonality", "back-transform", "jacobi/offsets", "Y = U'Z",
                                        "reconstruct+smooth"};
        unsigned long long ph[16];
        SHOPT_CUDA_TRY(cudaStreamSynchronize(st));
        SHOPT_CUDA_TRY(cudaMemcpyFromSymbol(ph, g_pca_phase, sizeof(ph)));
        fprintf(stderr, "[pca phases] stars %llu;", ph[0]);
        for (int i = 1; i < 12; i++) fprintf(stderr, " %s %.0f;", names[i], ph[0] ? (double)ph[i] / (double)ph[0] : 0.0);
        fprintf(stderr, " (cycles per star)\n");
    }
    return SHOPT_OK;
}
template <int RT>
static int pca_launch_nw(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    const int npairs = (a.c + 1) / 2;
    if (RT > 5) return pca_launch_t<16, RT>(a, scratch, scratch_bytes, st);
    // fewest (rounds x warps): 16, 22 or 32 warps
    const int cost16 = ((npairs + 15) / 16) * 16, cost22 = ((npairs + 21) / 22) * 22, cost32 = ((npairs + 31) / 32) * 32;
    if (RT <= 3 && cost32 < cost16 && cost32 < cost22) return pca_launch_t<32, RT>(a, scratch, scratch_bytes, st);
    if (cost22 < cost16) return pca_launch_t<22, RT>(a, scratch, scratch_bytes, st);
    return pca_launch_t<16, RT>(a, scratch, scratch_bytes, st);
}

// pca_kernel has ~20 instantiations (7 row-tile counts RT x up to 3 warp counts x 2 buffer placements), minutes of
// compile time each time this file changes.  The Makefile therefore compiles this file once per RT with -DSHOPT_PCA_RT=<RT>
// (objects pixelgrid_rt<RT>.o, built in parallel): such a translation unit emits only pca_launch_rt<RT>, the main one
// (no macro) everything else and no pca_kernel at all.
#ifdef SHOPT_PCA_RT
#define SHOPT_PCA_RT_FN2(rt) pca_launch_rt##rt
#define SHOPT_PCA_RT_FN(rt) SHOPT_PCA_RT_FN2(rt)
int SHOPT_PCA_RT_FN(SHOPT_PCA_RT)(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    return pca_launch_nw<SHOPT_PCA_RT>(a, scratch, scratch_bytes, st);
}
#else
int pca_launch_rt1(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);
int pca_launch_rt2(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);
int pca_launch_rt3(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);
int pca_launch_rt4(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);
int pca_launch_rt5(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);
int pca_launch_rt7(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);
int pca_launch_rt10(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st);

int pca_launch(const double *d_stamps, int r, int c, long long N, int ncomp, double pratio, int lanczos_a, int unity_sum,
               double *d_pgf, int *d_k, int *d_status, unsigned long long *d_counter, double *scratch,
               size_t scratch_bytes, cudaStream_t st) {
    if (!d_stamps || !d_pgf || !d_counter || r < 1 || c < 2 || N < 0 || ncomp < 1 || lanczos_a > 7) return SHOPT_E_BADARG;
    if (N == 0) return SHOPT_OK;
    if (r > 320 || c > 320) return SHOPT_E_UNSUPPORTED;
    PcaArgs a;
    a.stamps = d_stamps; a.r = r; a.c = c; a.N = N; a.ncomp = ncomp; a.pratio = pratio;
    a.lanczos_a = lanczos_a; a.unity_sum = unity_sum; a.pgf = d_pgf; a.k_out = d_k; a.status = d_status;
    a.counter = d_counter; a.wglobal = nullptr; a.wstride = 0;
    a.eigscratch = nullptr; a.eigstride = getenv("SHOPT_PCA_NO_EIG") ? 0 : (size_t)2 * r * r;
    a.kmax = ncomp < r ? ncomp : r; if (c < a.kmax) a.kmax = c;
    a.fast = (getenv("SHOPT_PCA_NO_FAST") == nullptr);
    a.debug = getenv("SHOPT_PCA_DEBUG") ? atoi(getenv("SHOPT_PCA_DEBUG")) : 0;   // 1: report iterations, 2: no L2 prefetch, 4: eig8 small-pair cut on, 8: first acceptance check after 2 iterations, 16: print cycles per phase, 32: no power path, 64: masked stars try the fast paths too
    if (lanczos_a > 0) lanczos_kernel2d(lanczos_a, a.kern, a.kz);
    a.list = nullptr; a.nlist = nullptr;
    // ---- stage 1: one-component stars by pca_power_kernel (pca_power.cu: no shared-memory copy of the stamp, several
    //      stars in flight per SM); what it does not accept lands on a work list that pca_kernel then serves ----
    const size_t lb = pca_list_bytes(N);
    if (a.fast && a.debug == 0 && !getenv("SHOPT_PCA_NO_LIGHT") && r >= 2 * FB && c >= 2 * FB && scratch &&
        scratch_bytes >= lb + pca_heavy_scratch_bytes(r, c, ncomp)) {
        PcaPowerArgs pa;
        pa.stamps = d_stamps; pa.r = r; pa.c = c; pa.N = N; pa.pratio = pratio; pa.lanczos_a = lanczos_a; pa.unity_sum = unity_sum;
        pa.pgf = d_pgf; pa.k_out = d_k; pa.status = d_status;
        pa.defer_list = reinterpret_cast<long long *>(scratch);
        pa.counter = reinterpret_cast<unsigned long long *>(pa.defer_list + N);
        pa.defer_count = pa.counter + 1;
        for (int i = 0; i < 15; i++) pa.kz[i] = lanczos_a > 0 ? a.kz[i] : 0.0;
        const int rcl = pca_power_launch(pa, st);
        if (rcl == SHOPT_OK) {
            a.list = pa.defer_list; a.nlist = pa.defer_count;
            scratch = reinterpret_cast<double *>(reinterpret_cast<char *>(scratch) + lb);
            scratch_bytes -= lb;
        } else if (rcl != SHOPT_E_UNSUPPORTED) {
            return rcl;
        }
    }
    const int rt = ((r > c ? r : c) + 31) / 32;
    switch (rt) {
        case 1: return pca_launch_rt1(a, scratch, scratch_bytes, st);
        case 2: return pca_launch_rt2(a, scratch, scratch_bytes, st);
        case 3: return pca_launch_rt3(a, scratch, scratch_bytes, st);
        case 4: return pca_launch_rt4(a, scratch, scratch_bytes, st);
        case 5: return pca_launch_rt5(a, scratch, scratch_bytes, st);
        case 6: case 7: return pca_launch_rt7(a, scratch, scratch_bytes, st);
        default: return pca_launch_rt10(a, scratch, scratch_bytes, st);
    }
}

int smooth_launch(const double *d_in, int r, int c, long long N, int lanczos_a, int nan_to_zero_flag, double *d_out,
                  cudaStream_t st) {
    if (!d_in || !d_out || r < 1 || c < 1 || N < 0 || lanczos_a > 7) return SHOPT_E_BADARG;
    if (N == 0) return SHOPT_OK;
    SmoothArgs a;
    a.in = d_in; a.out = d_out; a.N = N; a.r = r; a.c = c; a.la = lanczos_a; a.nan_to_zero = nan_to_zero_flag;
    for (int i = 0; i < 225; i++) a.kern[i] = 0.0;
    for (int i = 0; i < 15; i++) a.kz[i] = 0.0;
    if (lanczos_a > 0) lanczos_kernel2d(lanczos_a, a.kern, a.kz);
    const size_t fixed = (225 + 15) * sizeof(double), stamp = (size_t)r * c * sizeof(double);
    a.use_smem = (fixed + stamp <= 227 * 1024) ? 1 : 0;
    const size_t smem = fixed + (a.use_smem ? stamp : 0);
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(smooth_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    SHOPT_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, smooth_kernel, 512, smem));
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sm_count() * per_sm;
    if (grid > N) grid = N;
    smooth_kernel<<<(unsigned)grid, 512, smem, st>>>(a);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
#endif  // SHOPT_PCA_RT

}  // namespace shopt

// pca_eig.cuh -- direct path of pca_kernel for stars the block subspace iteration gives up on (k > 6: masked or noisy
// stamps whose spectrum is flat).  Instead of diagonalising all c columns by one-sided Jacobi (~5.6 ms per 131 x 131
// stamp) the r x r Gram matrix C = Z Z' is reduced to tridiagonal form by Householder reflectors, the leading
// eigenvalues are bracketed by Sturm-count multisection (8 lanes per eigenvalue), their eigenvectors come from inverse
// iteration on the tridiagonal matrix (pivoted LU, one thread per vector) and are carried back through the reflectors.
// Every step is checked a posteriori (tridiagonal residual, orthogonality, conditioning of the kept block); a star
// that fails any check falls through to the Jacobi SVD, so the results keep the accuracy the parity tests ask for.
// PCA semantics as in pixelgrid.cu (variances lambda_i / (c - 1), k by pratio, capped by ncomp).
#pragma once
#include <cfloat>

namespace shopt {

__device__ __forceinline__ double warp_sum_all(double v) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// C = Z Z' (r x r, full storage) into Cg.  Thread = (row i, block of four columns j); the i are consecutive over the
// lanes, so Z[i, t] is conflict-free and Z[j, t] a broadcast.
template <int NW>
__device__ void gram_rows(const double *Z, int r, int c, double *Cg) {
    const int nthr = NW * 32, nb = (r + 3) >> 2;
    for (int e = threadIdx.x; e < r * nb; e += nthr) {
        const int jb = e / r, i = e - jb * r, j0 = 4 * jb;
        if (j0 > i) continue;
        const int j1 = j0 + 1 < r ? j0 + 1 : r - 1, j2 = j0 + 2 < r ? j0 + 2 : r - 1, j3 = j0 + 3 < r ? j0 + 3 : r - 1;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        for (int t = 0; t < c; t++) {
            const double *z = Z + (size_t)t * r;
            const double zi = z[i];
            s0 = fma(zi, z[j0], s0);
            s1 = fma(zi, z[j1], s1);
            s2 = fma(zi, z[j2], s2);
            s3 = fma(zi, z[j3], s3);
        }
        Cg[i + (size_t)j0 * r] = s0; Cg[j0 + (size_t)i * r] = s0;
        Cg[i + (size_t)j1 * r] = s1; Cg[j1 + (size_t)i * r] = s1;
        Cg[i + (size_t)j2 * r] = s2; Cg[j2 + (size_t)i * r] = s2;
        Cg[i + (size_t)j3 * r] = s3; Cg[j3 + (size_t)i * r] = s3;
    }
}

// 1/q to ~1 ulp for normal q: the 20-bit hardware seed and two Newton steps (the Sturm recurrence and the reflector
// scaling need no IEEE division: a relative error of a few ulp is a relative perturbation of the same size)
__device__ __forceinline__ double rcp_newton(double q) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(q));
    r = fma(fma(-q, r, 1.0), r, r);
    r = fma(fma(-q, r, 1.0), r, r);
    return r;
}

// Householder reduction of the symmetric A (n x n, ld n, full storage, shared memory) to tridiagonal (d, e).  Reflector
// k is H_k = I - tau[k] u u' with u stored in Hg[k+1.., k] (global); A = H_0 ... H_{n-3} T H_{n-3} ... H_0.
// wk: 4 n doubles of shared scratch.  n <= 32 RT.
//
// One barrier and ONE pass over the trailing block per reflector:
//   * the rank-2 update of step k-1 is DELAYED: A holds A_old and (u, q) of the previous step stay in registers
//     (entry i on lane i & 31, register i >> 5, identical in every warp); whoever reads A applies  - (u_i q_j + q_i u_j)
//     on the fly, in the symmetric rounded form, so (i, j) and (j, i) keep identical bits;
//   * every warp forms column k and its reflector redundantly (same operations, same bits: no broadcast, no barrier);
//   * warp w owns the columns j = k+1 + w, k+1 + w + NW, ...: it reads column j once, applies the pending update, writes
//     it back and, the matrix being symmetric, dots it with the new u: that is row j of p = A u.  One warp reduction per
//     column, no cross-warp partial sums;
//   * after the step's only barrier every warp computes u'p and q = beta p - (beta^2 u'p / 2) u from the shared p.
// u and p are ping-pong buffered so that a fast warp's writes of step k+1 cannot overtake a slow warp's reads of step k.
// Measured on 131 x 131 (cycles per matrix, one CTA of 22 warps): 920k for the first version (four barrier-separated
// phases per step: reflector in warp 0, partial products, q in warp 0, rank-2 update), 801k for this one.  Two further
// restructurings were measured and dropped: the scalar work confined to warp 0 with a four-column pass over relative
// row indices (872k), and that plus u'p folded into the pass, rsqrt/rcp instead of sqrt and division (941k).  Thread
// 0's clock shows where a step's ~6k cycles go: ~2.4k in the reflector's scalar chain, ~3.6k in the pass, 30 cycles of
// barrier wait -- the routine is bound by the latency of dependent FP64 operations along 129 sequential steps, which
// only more independent work per SM (a second matrix; shared memory holds one) would hide.
template <int NW, int RT>
__device__ void tridiagonalise(double *A, int n, double *d, double *e, double *tau, double *wk, double *Hg) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    // (buffers addressed as offsets from wk, never through an array of pointers: the compiler then keeps the address
    //  space and emits LDS/STS instead of generic loads)
    for (int i = tid; i < 4 * n; i += nthr) wk[i] = 0.0;
    __syncthreads();
    double uprev[RT], qprev[RT];
#pragma unroll
    for (int t = 0; t < RT; t++) uprev[t] = qprev[t] = 0.0;
    double beta_prev = 0.0, K_prev = 0.0;
    // entry j of the pending q, from the shared p of the previous step (the formula the registers were filled with)
#define TRI_QPREV(j, upj) fma(-K_prev, (upj), beta_prev * pbp[(j)])
#define TRI_UPD(a, ui, qj, qi, uj) ((a) - __dadd_rn(__dmul_rn((ui), (qj)), __dmul_rn((qi), (uj))))
    for (int k = 0; k < n - 2; k++) {
        const int m0 = k + 1, cur = k & 1;
        const double *ubp = wk + (cur ^ 1) * n, *pbp = wk + (2 + (cur ^ 1)) * n;
        double *ubc = wk + cur * n, *pbc = wk + (2 + cur) * n;
        const double *colk = A + (size_t)k * n;
        // ---- column k of the current matrix and its reflector (every warp, redundantly) ----
        const double uk = ubp[k], qk = TRI_QPREV(k, uk);
        const double um = ubp[m0], qm = TRI_QPREV(m0, um);
        const double dk = TRI_UPD(colk[k], uk, qk, qk, uk);
        const double x0 = TRI_UPD(colk[m0], um, qk, qm, uk);
        double x[RT];
        double ss = 0.0;
#pragma unroll
        for (int t = 0; t < RT; t++) {
            const int i = lane + 32 * t;
            x[t] = 0.0;
            if (i > m0 && i < n) {
                x[t] = TRI_UPD(colk[i], uprev[t], qk, qprev[t], uk);
                ss = fma(x[t], x[t], ss);
            }
        }
        ss = warp_sum_all(ss);
        double alpha = x0, beta = 0.0, u0 = 0.0;
        if (ss > DBL_MIN) {
            const double nrm = sqrt(fma(x0, x0, ss));
            alpha = x0 > 0.0 ? -nrm : nrm;
            u0 = x0 - alpha;
            beta = 2.0 / fma(u0, u0, ss);
        }
        double ucur[RT];
#pragma unroll
        for (int t = 0; t < RT; t++) {
            const int i = lane + 32 * t;
            ucur[t] = (beta == 0.0 || i < m0 || i >= n) ? 0.0 : (i == m0 ? u0 : x[t]);
        }
        if (warp == 0) {
            if (lane == 0) { d[k] = dk; e[k] = alpha; tau[k] = beta; }
#pragma unroll
            for (int t = 0; t < RT; t++) {
                const int i = lane + 32 * t;
                if (i < n) ubc[i] = ucur[t];
                if (beta != 0.0 && i >= m0 && i < n) Hg[i + (size_t)k * n] = ucur[t];
            }
        }
        // ---- the pass: pending update applied to the owned columns, p_j = (column j) . u ----
        for (int j = m0 + warp; j < n; j += 2 * NW) {
            const int j2 = j + NW;
            const bool two = j2 < n;
            const double upj = ubp[j], qpj = TRI_QPREV(j, upj);
            const double upj2 = two ? ubp[j2] : 0.0, qpj2 = two ? TRI_QPREV(j2, upj2) : 0.0;
            double *cj = A + (size_t)j * n, *cj2 = A + (size_t)(two ? j2 : j) * n;
            double acc = 0.0, acc2 = 0.0;
#pragma unroll
            for (int t = 0; t < RT; t++) {
                const int i = lane + 32 * t;
                if (i >= m0 && i < n) {
                    const double a1 = TRI_UPD(cj[i], uprev[t], qpj, qprev[t], upj);
                    cj[i] = a1;
                    acc = fma(a1, ucur[t], acc);
                    if (two) {
                        const double a2 = TRI_UPD(cj2[i], uprev[t], qpj2, qprev[t], upj2);
                        cj2[i] = a2;
                        acc2 = fma(a2, ucur[t], acc2);
                    }
                }
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                acc += __shfl_xor_sync(0xffffffffu, acc, o);
                acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
            }
            if (lane == 0) { pbc[j] = acc; if (two) pbc[j2] = acc2; }
        }
        __syncthreads();
        // ---- q of this step (every warp, redundantly) ----
        double up = 0.0;
#pragma unroll
        for (int t = 0; t < RT; t++) {
            const int i = lane + 32 * t;
            qprev[t] = (i >= m0 && i < n) ? beta * pbc[i] : 0.0;       // beta p for now
            up = fma(ucur[t], qprev[t], up);
        }
        up = warp_sum_all(up);
        K_prev = 0.5 * beta * up;
        beta_prev = beta;
#pragma unroll
        for (int t = 0; t < RT; t++) {
            const int i = lane + 32 * t;
            qprev[t] = (i >= m0 && i < n) ? fma(-K_prev, ucur[t], qprev[t]) : 0.0;
            uprev[t] = ucur[t];
        }
    }
    if (tid == 0) {                                              // the last 2 x 2 block, pending update applied
        if (n >= 3) {
            const int last = (n - 3) & 1, a = n - 2, b = n - 1;
            const double *ubp = wk + last * n, *pbp = wk + (2 + last) * n;
            const double ua = ubp[a], ubb = ubp[b];
            const double qa = TRI_QPREV(a, ua), qb = TRI_QPREV(b, ubb);
            d[a] = TRI_UPD(A[a + (size_t)a * n], ua, qa, qa, ua);
            e[a] = TRI_UPD(A[b + (size_t)a * n], ubb, qa, qb, ua);
            d[b] = TRI_UPD(A[b + (size_t)b * n], ubb, qb, qb, ubb);
            tau[a] = 0.0;
        } else if (n == 2) {
            d[0] = A[0]; e[0] = A[1]; d[1] = A[3]; tau[0] = 0.0;
        } else {
            d[0] = A[0];
        }
        e[n - 1] = 0.0;
        tau[n - 1] = 0.0;
    }
#undef TRI_QPREV
#undef TRI_UPD
    __syncthreads();
}

// Number of eigenvalues of the tridiagonal (d, e) below x; e2[i] = e[i]^2.  Sturm sequence in its POLYNOMIAL form
//   p_0 = 1,  p_1 = d_0 - x,  p_{i+1} = (d_i - x) p_i - e_{i-1}^2 p_{i-1}:   count = sign changes along p_0 .. p_n,
// a zero taking the sign opposite to its predecessor's.  This loop runs on a few hundred threads whose only speed limit
// is the chain of DEPENDENT operations per step (346k of a masked star's 1.9M cycles went here): the quotient form
// q_i = p_i / p_{i-1} (LAPACK dstebz; the first version) has a reciprocal and ~6 dependent FP64 operations per step, the
// polynomial form ONE fused multiply-add -- d_i - x, e^2 p_{i-1}, the sign bookkeeping and the loads all hang off the
// chain and overlap once the loop is unrolled.  What the quotient form buys -- no overflow -- is recovered by rescaling
// (p_i, p_{i-1}) by a common power of two every eight steps: signs and the recurrence are invariant under it, and the
// caller hands over T scaled to |T| in [1, 2), so eight unscaled steps grow by at most 2^24 and can shrink by 2^-424.
// The loads are software-pipelined by hand (block i+1 is fetched into registers while block i's chain runs): left to
// itself the compiler issued each LDS right before its use, which put a shared-memory round trip on every step's chain.
#define STURM_STEP(dv, ev) do {                                          \
        const double c_ = fma((dv) - x, a, -((ev) * b));                 \
        const int sc_ = (c_ == 0.0) ? ~sa : __double2hiint(c_);          \
        cnt += (unsigned)(sc_ ^ sa) >> 31;                               \
        sa = sc_; b = a; a = c_;                                         \
    } while (0)
__device__ __forceinline__ int sturm_count(const double *d, const double *e2, int n, double x) {
    double b = 1.0, a = d[0] - x;                                // (p_{i-1}, p_i)
    int sa = (a == 0.0) ? (int)0x80000000 : __double2hiint(a);   // sign word of p_i under the zero rule
    int cnt = (unsigned)sa >> 31;
    int i = 1;
    if (n >= 9) {
        double d0 = d[1], d1 = d[2], d2 = d[3], d3 = d[4], f0 = e2[0], f1 = e2[1], f2 = e2[2], f3 = e2[3];
        for (; i + 11 < n; i += 8) {                             // steps i .. i+7; the block after them exists too
            const double g0 = d[i + 4], g1 = d[i + 5], g2 = d[i + 6], g3 = d[i + 7];
            const double h0 = e2[i + 3], h1 = e2[i + 4], h2 = e2[i + 5], h3 = e2[i + 6];
            STURM_STEP(d0, f0); STURM_STEP(d1, f1); STURM_STEP(d2, f2); STURM_STEP(d3, f3);
            d0 = d[i + 8]; d1 = d[i + 9]; d2 = d[i + 10]; d3 = d[i + 11];
            f0 = e2[i + 7]; f1 = e2[i + 8]; f2 = e2[i + 9]; f3 = e2[i + 10];
            STURM_STEP(g0, h0); STURM_STEP(g1, h1); STURM_STEP(g2, h2); STURM_STEP(g3, h3);
            // common power-of-two rescale: max(|a|, |b|) -> [1, 2)
            const int ea = __double2hiint(a) & 0x7ff00000, eb = __double2hiint(b) & 0x7ff00000;
            const double scl = __hiloint2double(0x7fe00000 - (ea > eb ? ea : eb), 0);
            a *= scl; b *= scl;
        }
        STURM_STEP(d0, f0); STURM_STEP(d1, f1); STURM_STEP(d2, f2); STURM_STEP(d3, f3);   // the block already in registers
        i += 4;
        const int ea = __double2hiint(a) & 0x7ff00000, eb = __double2hiint(b) & 0x7ff00000;
        const double scl = __hiloint2double(0x7fe00000 - (ea > eb ? ea : eb), 0);
        a *= scl; b *= scl;
    }
    for (; i < n; i++) STURM_STEP(d[i], e2[i - 1]);              // at most 11 steps
    return cnt;
}
#undef STURM_STEP

// lam[j], j < m: the m largest eigenvalues in descending order.  Eight lanes per eigenvalue split the bracket into nine
// parts per round (N(lo) < n - j <= N(hi) throughout).
template <int NW>
__device__ void top_eigenvalues(const double *d, const double *e2, int n, int m, double glo, double ghi, double pivmin,
                                double *lam) {
    // (pivmin only enters the stopping rule now)
    const int tid = threadIdx.x, lane = tid & 31, sub = tid & 7, ngrp = NW * 4, sh = lane & 24;
    const unsigned gmask = 0xffu << sh;
    for (int j = tid >> 3; j < m; j += ngrp) {
        const int want = n - j;
        double lo = glo, hi = ghi;
        for (int it = 0; it < 26; it++) {
            const double w = (hi - lo) * (1.0 / 9.0);
            const double x = fma(w, (double)(sub + 1), lo);
            const int cnt = sturm_count(d, e2, n, x);
            const unsigned b = (__ballot_sync(gmask, cnt >= want) >> sh) & 0xffu;
            const int s = b ? __ffs(b) - 1 : 8;
            const double nlo = s == 0 ? lo : fma(w, (double)s, lo);
            const double nhi = s == 8 ? hi : fma(w, (double)(s + 1), lo);
            lo = nlo; hi = nhi;
            if (hi - lo <= 2.0 * DBL_EPSILON * fmax(fabs(lo), fabs(hi)) + 2.0 * pivmin) break;
        }
        if (sub == 0) lam[j] = 0.5 * (lo + hi);
    }
}

// 1/x to ~1 ulp for normal x well inside the exponent range (pivots of T - lam I, floored at eps |T|)
__device__ __forceinline__ double rcp_pivot(double x) {
    return (fabs(x) > 1e-290 && fabs(x) < 1e290) ? rcp_newton(x) : 1.0 / x;
}

// Inverse iteration for one eigenvector of the tridiagonal (d, e) at the shift lamj: LU of T - lamj I with partial
// pivoting (multipliers F, reciprocal pivots RD, swap flags SW, all interleaved over the kst vectors of the CTA: entry
// i of vector j at [i kst + j]); the super-diagonals of U follow from the flags.  v (n doubles) receives the unit
// eigenvector; returns the squared residual |T v - lamj v|^2.
__device__ __forceinline__ double tri_inverse_iteration(const double *d, const double *e, int n, double lamj, int j, int kst, double *F,
                                        double *RD, int *SW, double *v, double tnorm) {
    const double tiny = fmax(DBL_EPSILON * tnorm, DBL_MIN);
    double di = d[0] - lamj, dui = e[0];
    for (int i = 0; i < n - 1; i++) {
        const double dli = e[i], dnext = d[i + 1] - lamj, dunext = e[i + 1];
        const size_t o = (size_t)i * kst + j;
        if (fabs(di) >= fabs(dli)) {
            if (di == 0.0) di = tiny;
            const double rd = rcp_pivot(di), f = dli * rd;
            F[o] = f; RD[o] = rd; SW[o] = 0;
            di = dnext - f * dui;
            dui = dunext;
        } else {
            const double rd = rcp_pivot(dli), f = di * rd;
            F[o] = f; RD[o] = rd; SW[o] = 1;
            di = dui - f * dnext;
            dui = -f * dunext;
        }
    }
    if (fabs(di) < tiny) di = di < 0.0 ? -tiny : tiny;
    RD[(size_t)(n - 1) * kst + j] = 1.0 / di;
    // start vector: fixed pseudo-random entries in (0.5, 1.5), different for every j
    for (int i = 0; i < n; i++) {
        unsigned h = (unsigned)(i + 1) * 2654435761u + (unsigned)(j + 1) * 2246822519u;
        h ^= h >> 15; h *= 2654435761u; h ^= h >> 13;
        v[i] = 0.5 + (double)(h >> 8) * (1.0 / 16777216.0);
    }
    double scale = 1.0;
    for (int it = 0; it < 3; it++) {
        if (it > 0) {                                            // forward: b <- L^-1 P b  (the first pass skips it)
            double bi = v[0] * scale;
            for (int i = 0; i < n - 1; i++) {
                const size_t o = (size_t)i * kst + j;
                const double bn = v[i + 1] * scale, f = F[o];
                if (SW[o]) { v[i] = bn; bi = bi - f * bn; }
                else { v[i] = bi; bi = bn - f * bi; }
            }
            v[n - 1] = bi;
        }
        double x1 = v[n - 1] * RD[(size_t)(n - 1) * kst + j], x2 = 0.0, mx = fabs(x1);
        v[n - 1] = x1;
        for (int i = n - 2; i >= 0; i--) {                       // backward: b <- U^-1 b
            const size_t o = (size_t)i * kst + j;
            const int sw = SW[o];
            double du, du2;
            if (sw) { du = d[i + 1] - lamj; du2 = e[i + 1]; }
            else { du = (i > 0 && SW[o - kst]) ? -F[o - kst] * e[i] : e[i]; du2 = 0.0; }
            const double x = (v[i] - du * x1 - du2 * x2) * RD[o];
            v[i] = x;
            mx = fmax(mx, fabs(x));
            x2 = x1; x1 = x;
        }
        if (!(mx > 0.0) || !isfinite(mx)) return INFINITY;
        scale = 1.0 / mx;
    }
    double nn = 0.0;
    for (int i = 0; i < n; i++) { const double x = v[i] * scale; nn = fma(x, x, nn); }
    const double sc = scale / sqrt(nn);
    for (int i = 0; i < n; i++) v[i] *= sc;
    double res = 0.0;
    for (int i = 0; i < n; i++) {
        double t = (d[i] - lamj) * v[i];
        if (i > 0) t = fma(e[i - 1], v[i - 1], t);
        if (i < n - 1) t = fma(e[i], v[i + 1], t);
        res = fma(t, t, res);
    }
    return res;
}

// Direct PCA of the centred stamp Z (r x c in W, r <= c, r <= NW*32).  On success returns k > 0 with the orthonormal
// U_k in V (column a at V + a r) and the variances' numerators in lam[0..k); W is free on return.  Returns 0 when a
// check fails (the caller reloads Z and runs the Jacobi SVD).
//   det: 3 r doubles (d, e, tau), region: the Y / fast-path area behind W (>= max(kmax, 10) r doubles), fth: 8 doubles,
//   Cg, Hg: r*r doubles each of per-CTA global scratch.
template <int NW, int RT>
__device__ int gram_eig_pca(double *W, int r, int c, int ncomp, double pratio, double *region, double *det, double *lam,
                            double *fth, int *s_flag, double *Cg, double *Hg, unsigned long long *prof) {
#define EIG_TICK(i) do { if (prof && threadIdx.x == 0) prof[i] = clock64(); } while (0)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32, n = r;
    double *d = det, *e = det + n, *tau = det + 2 * n;
    double *V = region;
    gram_rows<NW>(W, r, c, Cg);
    __syncthreads();
    for (int p = tid; p < n * n; p += nthr) W[p] = Cg[p];
    __syncthreads();
    EIG_TICK(3);
    tridiagonalise<NW, RT>(W, n, d, e, tau, region, Hg);
    EIG_TICK(4);
    // Gershgorin bracket, trace, norm
    double *e2 = region;
    if (warp == 0) {
        double lo = INFINITY, hi = -INFINITY, tr = 0.0;
        for (int i = lane; i < n; i += 32) {
            const double rad = fabs(e[i]) + (i > 0 ? fabs(e[i - 1]) : 0.0);
            lo = fmin(lo, d[i] - rad);
            hi = fmax(hi, d[i] + rad);
            tr += d[i];
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            tr += __shfl_xor_sync(0xffffffffu, tr, o);
        }
        if (lane == 0) {
            const double tn = fmax(fabs(lo), fabs(hi));
            fth[0] = lo - 2.0 * DBL_EPSILON * n * tn - DBL_MIN;
            fth[1] = hi + 2.0 * DBL_EPSILON * n * tn + DBL_MIN;
            fth[2] = tr;
            fth[3] = tn;
        }
    }
    __syncthreads();
    const double tnorm = fth[3];
    if (!(tnorm > 1e-200 && tnorm < 1e200)) return 0;            // zero, non-finite or absurd: let the Jacobi SVD decide
    // the eigenvalue search runs on T scaled by an exact power of two to |T| in [1, 2) (sturm_count's contract)
    const int texp = __double2hiint(tnorm) & 0x7ff00000;
    const double tsc = __hiloint2double(0x7fe00000 - texp, 0), tinv = __hiloint2double(texp, 0);
    double *ds = region + n;
    for (int i = tid; i < n; i += nthr) { const double es = e[i] * tsc; e2[i] = es * es; ds[i] = d[i] * tsc; }
    __syncthreads();
    int m = ncomp < n ? ncomp : n;
    if (c < m) m = c;
    top_eigenvalues<NW>(ds, e2, n, m, fth[0] * tsc, fth[1] * tsc, 4.0 * DBL_MIN, lam);
    __syncthreads();
    for (int i = tid; i < m; i += nthr) lam[i] *= tinv;
    __syncthreads();
    EIG_TICK(5);
    if (tid == 0) {                                              // choose_pcadim on lambda_i / (c - 1)
        const double den = (double)(c - 1);
        const double thres = (fth[2] / den) * pratio;
        int kk = 1;
        double acc = lam[0] / den;
        while (kk < m && acc < thres) { acc += lam[kk] / den; kk++; }
        // the Gram matrix squares the condition number: keep this path to well-conditioned kept blocks
        if (!(lam[0] > 0.0) || !(lam[kk - 1] > 1e-6 * lam[0]) || 5 * kk > 2 * c) kk = 0;
        *s_flag = kk;
    }
    __syncthreads();
    const int k = *s_flag;
    if (k == 0) return 0;
    // ---- eigenvectors of T: inverse iteration, one thread per vector, workspace in W ----
    int bad = 0;
    if (tid < k) {
        const int j = tid;
        const double sep = 10.0 * DBL_EPSILON * tnorm;
        int cl = 0;
        for (int i = j - 1; i >= 0 && lam[i] - lam[i + 1] < sep; i--) cl++;
        const double lamj = lam[j] - cl * sep;
        const double res = tri_inverse_iteration(d, e, n, lamj, j, k, W, W + (size_t)n * k,
                                                 reinterpret_cast<int *>(W + (size_t)2 * n * k), V + (size_t)j * n, tnorm);
        const double tol = 1e-12 * tnorm;
        bad = !(res <= tol * tol);
    }
    bad = __syncthreads_or(bad);
    EIG_TICK(6);
    if (bad) return 0;
    // ---- orthogonality of the k vectors; tight clusters are re-orthogonalised (MGS, twice) by warp 0 ----
    int loose = 0;
    for (int pr = warp; pr < k * k; pr += NW) {
        const int i = pr / k, j = pr - i * k;
        if (i >= j) continue;
        double s = 0.0;
        for (int t = lane; t < n; t += 32) s = fma(V[(size_t)i * n + t], V[(size_t)j * n + t], s);
        s = warp_sum_all(s);
        if (fabs(s) > 1e-12) loose = 1;
    }
    loose = __syncthreads_or(loose);
    if (loose) {
        if (warp == 0) {
            for (int j = 0; j < k; j++) {
                double vj[RT];
#pragma unroll
                for (int t = 0; t < RT; t++) vj[t] = (lane + 32 * t < n) ? V[(size_t)j * n + lane + 32 * t] : 0.0;
                for (int pass = 0; pass < 2; pass++)
                    for (int i = 0; i < j; i++) {
                        double s = 0.0;
#pragma unroll
                        for (int t = 0; t < RT; t++)
                            if (lane + 32 * t < n) s = fma(V[(size_t)i * n + lane + 32 * t], vj[t], s);
                        s = warp_sum_all(s);
#pragma unroll
                        for (int t = 0; t < RT; t++)
                            if (lane + 32 * t < n) vj[t] = fma(-s, V[(size_t)i * n + lane + 32 * t], vj[t]);
                    }
                double nn = 0.0;
#pragma unroll
                for (int t = 0; t < RT; t++) nn = fma(vj[t], vj[t], nn);
                nn = warp_sum_all(nn);
                if (!(nn > 0.25)) bad = 1;                       // the vector was (nearly) inside the span of the others
                const double sc = rsqrt(nn);
#pragma unroll
                for (int t = 0; t < RT; t++)
                    if (lane + 32 * t < n) V[(size_t)j * n + lane + 32 * t] = vj[t] * sc;
                __syncwarp();
            }
        }
        bad = __syncthreads_or(bad);
        if (bad) return 0;
    }
    EIG_TICK(7);
    // ---- back through the reflectors: u_a = H_0 ... H_{n-3} v_a, one warp per vector ----
    for (int p = tid; p < n * (n - 2); p += nthr) {
        const int kk = p / n, i = p - kk * n;
        if (i > kk) W[p] = Hg[p];
    }
    __syncthreads();
    for (int a = warp; a < k; a += NW) {
        double va[RT];
#pragma unroll
        for (int t = 0; t < RT; t++) va[t] = (lane + 32 * t < n) ? V[(size_t)a * n + lane + 32 * t] : 0.0;
        for (int kk = n - 3; kk >= 0; kk--) {
            const double beta = tau[kk];
            if (beta == 0.0) continue;
            const double *uk = W + (size_t)kk * n;
            double uu[RT];
            double s = 0.0;
#pragma unroll
            for (int t = 0; t < RT; t++) {
                const int i = lane + 32 * t;
                uu[t] = (i > kk && i < n) ? uk[i] : 0.0;
                s = fma(uu[t], va[t], s);
            }
            s = warp_sum_all(s) * beta;
#pragma unroll
            for (int t = 0; t < RT; t++) va[t] = fma(-s, uu[t], va[t]);
        }
#pragma unroll
        for (int t = 0; t < RT; t++)
            if (lane + 32 * t < n) V[(size_t)a * n + lane + 32 * t] = va[t];
    }
    __syncthreads();
    EIG_TICK(8);
    return k;
#undef EIG_TICK
}

}  // namespace shopt

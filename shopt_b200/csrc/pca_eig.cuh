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

// Householder reduction of the symmetric A (n x n, ld n, full storage, shared memory) to tridiagonal (d, e).  Reflector
// k is H_k = I - tau[k] u u' with u stored in Hg[k+1.., k] (global); A = H_0 ... H_{n-3} T H_{n-3} ... H_0.
// u, q: n doubles each, pp: NP n doubles (shared scratch).  n <= NW*32.
template <int NW, int RT>
__device__ void tridiagonalise(double *A, int n, double *d, double *e, double *tau, double *u, double *q, double *pp,
                               double *Hg) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    int NP = nthr / n;
    if (NP > 8) NP = 8;
    const int prow = tid % n, ppart = tid / n;
    for (int k = 0; k < n - 2; k++) {
        const int m0 = k + 1, len = n - m0;
        const double *xk = A + (size_t)k * n;
        if (warp == 0) {
            double ss = 0.0;
            for (int i = m0 + 1 + lane; i < n; i += 32) ss = fma(xk[i], xk[i], ss);
            ss = warp_sum_all(ss);
            const double x0 = xk[m0];
            double alpha = x0, beta = 0.0, u0 = 0.0;
            if (ss > DBL_MIN) {
                const double nrm = sqrt(fma(x0, x0, ss));
                alpha = x0 > 0.0 ? -nrm : nrm;
                u0 = x0 - alpha;
                beta = 2.0 / fma(u0, u0, ss);
            }
            for (int i = lane; i < n; i += 32) u[i] = (i < m0 || beta == 0.0) ? 0.0 : (i == m0 ? u0 : xk[i]);
            if (lane == 0) { d[k] = xk[k]; e[k] = alpha; tau[k] = beta; }
        }
        __syncthreads();
        const double beta = tau[k];
        if (beta != 0.0) {                                       // uniform over the CTA
            if (ppart < NP && prow >= m0) {                      // partial products of A22 u
                const int seg = (len + NP - 1) / NP, j0 = m0 + ppart * seg, j1 = j0 + seg < n ? j0 + seg : n;
                double s = 0.0;
                for (int j = j0; j < j1; j++) s = fma(A[prow + (size_t)j * n], u[j], s);
                pp[ppart * n + prow] = s;
            }
            __syncthreads();
            if (warp == 0) {                                     // p = beta A22 u,  q = p - (beta u'p / 2) u
                double pv[RT];
                double up = 0.0;
#pragma unroll
                for (int t = 0; t < RT; t++) {
                    const int i = lane + 32 * t;
                    pv[t] = 0.0;
                    if (i >= m0 && i < n) {
                        double x[8];                             // NP <= 8 partial sums, loaded together
#pragma unroll
                        for (int part = 0; part < 8; part++) x[part] = part < NP ? pp[part * n + i] : 0.0;
                        const double s = ((x[0] + x[1]) + (x[2] + x[3])) + ((x[4] + x[5]) + (x[6] + x[7]));
                        pv[t] = beta * s;
                        up = fma(u[i], pv[t], up);
                    }
                }
                up = warp_sum_all(up);
                const double K = 0.5 * beta * up;
#pragma unroll
                for (int t = 0; t < RT; t++) {
                    const int i = lane + 32 * t;
                    if (i < n) q[i] = (i >= m0) ? pv[t] - K * u[i] : 0.0;
                }
            }
            __syncthreads();
            {                                                    // A22 -= u q' + q u'   (both triangles)
                const int row = tid % len, col0 = tid / len, cstep = nthr / len;
                if (col0 < cstep) {
                    const int i = m0 + row;
                    const double ui = u[i], qi = q[i];
                    for (int j = m0 + col0; j < n; j += cstep) {
                        const double t2 = __dadd_rn(__dmul_rn(ui, q[j]), __dmul_rn(qi, u[j]));
                        A[i + (size_t)j * n] -= t2;
                    }
                }
            }
            for (int i = m0 + tid; i < n; i += nthr) Hg[i + (size_t)k * n] = u[i];
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (n >= 2) {
            d[n - 2] = A[(n - 2) + (size_t)(n - 2) * n];
            e[n - 2] = A[(n - 1) + (size_t)(n - 2) * n];
            tau[n - 2] = 0.0;
        }
        d[n - 1] = A[(n - 1) + (size_t)(n - 1) * n];
        e[n - 1] = 0.0;
        tau[n - 1] = 0.0;
    }
    __syncthreads();
}

// 1/q to ~1 ulp for normal q: the 20-bit hardware seed and two Newton steps (the Sturm recurrence needs no IEEE division:
// a relative error of a few ulp is a relative perturbation of e^2 of the same size)
__device__ __forceinline__ double rcp_newton(double q) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(q));
    r = fma(fma(-q, r, 1.0), r, r);
    r = fma(fma(-q, r, 1.0), r, r);
    return r;
}

// number of eigenvalues of the tridiagonal (d, e) below x; e2[i] = e[i]^2; pivmin is a normal number
__device__ __forceinline__ int sturm_count(const double *d, const double *e2, int n, double x, double pivmin) {
    double q = d[0] - x;
    if (fabs(q) < pivmin) q = -pivmin;
    int cnt = q < 0.0;
    for (int i = 1; i < n; i++) {
        q = fma(-e2[i - 1], rcp_newton(q), d[i] - x);
        if (fabs(q) < pivmin) q = -pivmin;
        cnt += q < 0.0;
    }
    return cnt;
}

// lam[j], j < m: the m largest eigenvalues in descending order.  Eight lanes per eigenvalue split the bracket into nine
// parts per round (N(lo) < n - j <= N(hi) throughout).
template <int NW>
__device__ void top_eigenvalues(const double *d, const double *e2, int n, int m, double glo, double ghi, double pivmin,
                                double *lam) {
    const int tid = threadIdx.x, lane = tid & 31, sub = tid & 7, ngrp = NW * 4, sh = lane & 24;
    const unsigned gmask = 0xffu << sh;
    for (int j = tid >> 3; j < m; j += ngrp) {
        const int want = n - j;
        double lo = glo, hi = ghi;
        for (int it = 0; it < 26; it++) {
            const double w = (hi - lo) * (1.0 / 9.0);
            const double x = fma(w, (double)(sub + 1), lo);
            const int cnt = sturm_count(d, e2, n, x, pivmin);
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
    tridiagonalise<NW, RT>(W, n, d, e, tau, region, region + n, region + 2 * n, Hg);
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
    for (int i = tid; i < n; i += nthr) e2[i] = e[i] * e[i];
    __syncthreads();
    const double tnorm = fth[3];
    int m = ncomp < n ? ncomp : n;
    if (c < m) m = c;
    top_eigenvalues<NW>(d, e2, n, m, fth[0], fth[1], DBL_MIN * fmax(1.0, tnorm * tnorm), lam);
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

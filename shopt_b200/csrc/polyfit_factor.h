// polyfit_factor.h -- the K x K part of polynomial_optimizer's  A \ z  (interpolate.jl:131-138), shared by the device
// kernel (polyfit.cu: factor_kernel) and a host entry point (shopt_polyfit_factor_host, CPU-testable).
//
// The reference builds A[i,k] = u_i^p v_i^q from RAW coordinates and solves by LAPACK's column-pivoted QR with rank
// truncation (rcond = min(n,K) eps) and minimum-norm completion.  With sky coordinates (u ~ 150, v ~ 2, field 0.2 deg)
// cond(A) is 1e11..1e17, so V'V in that basis is useless.  Here the streaming pass accumulates the normal equations in
// a CENTRED and SCALED basis  ut = (u - u0)/su, vt = (v - v0)/sv  (cond(Vc) ~ 10..100):  Gc = Vc'Vc,  b = Vc'z.
// Since  A = Vc T  with the K x K change of basis T (binomial expansion of (su ut + u0)^i (sv vt + v0)^j):
//     Gc[pi,pi] = L L'  (equilibrated, diagonally pivoted Cholesky, rank rc)   =>   Vc = Qc Rc,  Rc = L' Pi' D^-1
//     A = Qc (Rc T);   M = Rc T  (rc x K) is factored by column-pivoted Householder QR:  M Pm = Qm Rm
// which is the pivoted QR of A itself (Qc has orthonormal columns, so column norms, pivot order and the diagonal of
// Rm are those LAPACK sees), with the same rank rule.  Per pixel only  y = Qc'z = L1^-1 (D b)[pi]  is needed:
//     full rank      c_raw = S D Pi L^-T y               (S: exact binomial map centred -> raw coefficients)
//     rank < K       c_raw = Pm R1'(R1 R1')^-1 Qm1' y     (minimum norm, like LAPACK's complete orthogonal factorisation)
// Both are ONE K x rc matrix W2 applied after a forward substitution: the solve kernel is the same in every case.
#pragma once
#include <math.h>

#include "select_core.h"   // SHOPT_HD

namespace shopt {

constexpr int KMAX = 45;   // degree 8

#define SHOPT_POLY_RANK_DEFICIENT 1   /* raw design matrix numerically rank deficient: minimum-norm answer, like A \ z */
#define SHOPT_POLY_NONFINITE      2   /* NaN / Inf in the normal equations: coefficients are NaN                      */
#define SHOPT_POLY_ILL_CONDITIONED 4  /* centred design nearly dependent (pivot < 1e-8): normal equations lose digits  */

struct PolyFactor {
    int K, rc_rank, rank, info;
    int piv[KMAX];
    int jp[KMAX];             // column permutation of the change-of-basis QR (work space of polyfit_factor)
    double dsc[KMAX];
    double L[KMAX * KMAX];    // [m][k], m in pivoted order
    double W2[KMAX * KMAX];   // [j][i]: c_raw[j] = sum_{i < rc_rank} W2[j][i] y[i]
};

SHOPT_HD double binom(int n, int k) {
    double r = 1.0;
    for (int i = 1; i <= k; i++) r = r * (double)(n - k + i) / (double)i;
    return r;
}
SHOPT_HD double ipow_(double x, int n) { double r = 1.0; for (int i = 0; i < n; i++) r *= x; return r; }

// work: 7 K^2 + 2 K doubles.  G: K x K row-major.  ctr = {u0, 1/su, v0, 1/sv} or null (centred basis == raw basis).
// nrows: rows of the design matrix (training stars).  drop_const: solve without the raw constant column (its
// coefficient comes out 0) -- median_constraint with a zero first column (interpolate.jl:118-124).
//
// lane / nl: the routine is written for `nl` cooperating lanes of ONE warp (device: factor_kernel, nl = 32, work in
// shared memory) or a single caller (host: lane 0 of 1).  Every matrix entry is computed by one lane with the serial
// code's operations in the serial code's order, so the result does not depend on nl; the lanes only share out the
// entries and meet at FACTOR_SYNC between phases.  On one thread the K = 21 factorisation took 2.9 ms (12 % of a
// degree-5 stage step: K^2 entries of the two changes of basis each re-deriving binomials and powers inside a K-term
// sum); shared out, with the basis maps formed once, it is tens of microseconds.
#ifdef __CUDA_ARCH__
#define FACTOR_SYNC() __syncwarp()
#else
#define FACTOR_SYNC() ((void)0)
#endif
SHOPT_HD void polyfit_factor(int degree, const double *G, const double *ctr, long long nrows, int drop_const, double *work,
                             PolyFactor *F, int lane = 0, int nl = 1) {
    const int K = (degree + 1) * (degree + 2) / 2;
    double *A = work, *T = work + K * K, *M = work + 2 * K * K, *Qm = work + 3 * K * K, *Li = work + 4 * K * K,
           *Cc = work + 5 * K * K, *S = work + 6 * K * K, *cn = work + 7 * K * K, *wv = work + 7 * K * K + K;
    int tp[KMAX], tq[KMAX];
    {
        int k = 0;
        for (int p = 0; p <= degree; p++) for (int q = 0; q <= degree - p; q++) { tp[k] = p; tq[k] = q; k++; }
    }
    double *L = F->L, *W2 = F->W2, *d = F->dsc;
    int *jp = F->jp;
    if (lane == 0) { F->K = K; F->info = 0; }
    FACTOR_SYNC();
    int bad = 0;
    for (int i = lane; i < K * K; i += nl) { L[i] = 0.0; W2[i] = 0.0; if (!isfinite(G[i])) bad = 1; }
    if (ctr) for (int i = 0; i < 4; i++) if (!isfinite(ctr[i])) bad = 1;
#ifdef __CUDA_ARCH__
    bad = __any_sync(0xffffffffu, bad);
#endif
    if (bad) {
        FACTOR_SYNC();
        if (lane == 0) {
            F->info |= SHOPT_POLY_NONFINITE;
            for (int i = 0; i < K; i++) { F->piv[i] = i; d[i] = 1.0; L[i * K + i] = 1.0; }
            for (int i = 0; i < K * K; i++) W2[i] = nan("");
            F->rc_rank = K; F->rank = 0;
        }
        FACTOR_SYNC();
        return;
    }
    // 1. equilibrate
    for (int i = lane; i < K; i += nl) { d[i] = (G[i * K + i] > 0.0) ? 1.0 / sqrt(G[i * K + i]) : 0.0; F->piv[i] = i; }
    FACTOR_SYNC();
    for (int e = lane; e < K * K; e += nl) { const int i = e / K, j = e % K; A[e] = G[e] * d[i] * d[j]; }
    FACTOR_SYNC();
    // 2. diagonally pivoted Cholesky  A[piv,piv] = L L'
    int rc = 0;
    double minpiv = 1.0;
    for (int k = 0; k < K; k++) {
        int pv = k;                                       // (every lane finds the same pivot: same reads, same compares)
        for (int i = k + 1; i < K; i++) if (A[F->piv[i] * K + F->piv[i]] > A[F->piv[pv] * K + F->piv[pv]]) pv = i;
        FACTOR_SYNC();
        if (lane == 0) { const int t = F->piv[k]; F->piv[k] = F->piv[pv]; F->piv[pv] = t; }
        for (int c = lane; c < k; c += nl) { const double t = L[k * K + c]; L[k * K + c] = L[pv * K + c]; L[pv * K + c] = t; }
        FACTOR_SYNC();
        const int p = F->piv[k];
        const double a = A[p * K + p];
        if (!(a > 1e-13)) break;
        if (a < minpiv) minpiv = a;
        const double l = sqrt(a);
        if (lane == 0) L[k * K + k] = l;
        for (int i = k + 1 + lane; i < K; i += nl) L[i * K + k] = A[F->piv[i] * K + p] / l;
        FACTOR_SYNC();
        const int mrem = K - k - 1;
        for (int e = lane; e < mrem * mrem; e += nl) {
            const int i = k + 1 + e / mrem, j = k + 1 + e % mrem;
            if (j <= i) {
                const double v = A[F->piv[i] * K + F->piv[j]] - L[i * K + k] * L[j * K + k];
                A[F->piv[i] * K + F->piv[j]] = v; A[F->piv[j] * K + F->piv[i]] = v;
            }
        }
        FACTOR_SYNC();
        rc++;
    }
    if (lane == 0) {
        F->rc_rank = rc;
        if (minpiv < 1e-8) F->info |= SHOPT_POLY_ILL_CONDITIONED;
    }
    for (int i = rc + lane; i < K; i += nl) L[i * K + i] = 1.0;     // keeps the (unused) tail of the forward substitution finite
    // 3. T: raw monomial kr = (i,j) in centred monomials kc = (p,q);  S: its inverse (centred -> raw), used in 6a
    const double u0 = ctr ? ctr[0] : 0.0, su = ctr ? 1.0 / ctr[1] : 1.0, v0 = ctr ? ctr[2] : 0.0, sv = ctr ? 1.0 / ctr[3] : 1.0;
    {
        const double ia = 1.0 / su, ib = -u0 / su, ic = 1.0 / sv, ie = -v0 / sv;
        for (int e = lane; e < K * K; e += nl) {
            const int kc = e / K, kr = e % K;
            const int p = tp[kc], q = tq[kc], i = tp[kr], j = tq[kr];
            double t = 0.0;
            if (p <= i && q <= j) t = binom(i, p) * ipow_(su, p) * ipow_(u0, i - p) * binom(j, q) * ipow_(sv, q) * ipow_(v0, j - q);
            if (drop_const && kr == 0) t = 0.0;
            T[kc * K + kr] = t;
            // S[kr][kc]: coefficient of raw monomial kr = (i, j) in centred monomial kc = (p, q)
            double sv_ = 0.0;
            if (i <= p && j <= q) sv_ = binom(p, i) * ipow_(ia, i) * ipow_(ib, p - i) * binom(q, j) * ipow_(ic, j) * ipow_(ie, q - j);
            S[kr * K + kc] = sv_;
        }
    }
    FACTOR_SYNC();
    // 4. M = Rc T,  Rc[i][piv[m]] = L[m][i] / d[piv[m]]
    for (int e = lane; e < rc * K; e += nl) {
        const int i = e / K, kr = e % K;
        double s = 0.0;
        for (int m = i; m < K; m++) {
            const int c = F->piv[m];
            if (d[c] != 0.0) s += (L[m * K + i] / d[c]) * T[c * K + kr];
        }
        M[i * K + kr] = s;
    }
    // 5. column-pivoted Householder QR of M (rc x K); Qm (rc x rc) accumulated explicitly
    for (int e = lane; e < rc * rc; e += nl) { const int i = e / rc, j = e % rc; Qm[i * K + j] = (i == j) ? 1.0 : 0.0; }
    for (int j = lane; j < K; j += nl) jp[j] = j;
    FACTOR_SYNC();
    const int mn = rc < K ? rc : K;
    for (int k = 0; k < mn; k++) {
        for (int j = k + lane; j < K; j += nl) {          // squared norms of the trailing columns
            double s = 0.0;
            for (int i = k; i < rc; i++) s += M[i * K + j] * M[i * K + j];
            cn[j] = s;
        }
        FACTOR_SYNC();
        int pv = k; double best = -1.0;
        for (int j = k; j < K; j++) if (cn[j] > best) { best = cn[j]; pv = j; }
        FACTOR_SYNC();
        if (pv != k) {
            for (int i = lane; i < rc; i += nl) { const double t = M[i * K + k]; M[i * K + k] = M[i * K + pv]; M[i * K + pv] = t; }
            if (lane == 0) { const int t = jp[k]; jp[k] = jp[pv]; jp[pv] = t; }
        }
        FACTOR_SYNC();
        double xn = 0.0;
        for (int i = k + 1; i < rc; i++) xn += M[i * K + k] * M[i * K + k];
        const double alpha = M[k * K + k];
        if (xn == 0.0) continue;                          // H = I   (uniform: every lane read the same column)
        const double beta = -copysign(sqrt(alpha * alpha + xn), alpha);
        const double tau = (beta - alpha) / beta, sc = 1.0 / (alpha - beta);
        FACTOR_SYNC();
        for (int i = k + 1 + lane; i < rc; i += nl) M[i * K + k] *= sc;   // v (v[k] = 1 implicit)
        if (lane == 0) M[k * K + k] = beta;
        FACTOR_SYNC();
        for (int j = k + 1 + lane; j < K; j += nl) {
            double w = M[k * K + j];
            for (int i = k + 1; i < rc; i++) w += M[i * K + k] * M[i * K + j];
            w *= tau;
            M[k * K + j] -= w;
            for (int i = k + 1; i < rc; i++) M[i * K + j] -= w * M[i * K + k];
        }
        for (int r = lane; r < rc; r += nl) {             // Qm <- Qm H
            double w = Qm[r * K + k];
            for (int i = k + 1; i < rc; i++) w += Qm[r * K + i] * M[i * K + k];
            w *= tau;
            Qm[r * K + k] -= w;
            for (int i = k + 1; i < rc; i++) Qm[r * K + i] -= w * M[i * K + k];
        }
        FACTOR_SYNC();
        for (int i = k + 1 + lane; i < rc; i += nl) M[i * K + k] = 0.0;
        FACTOR_SYNC();
    }
    // numerical rank, the oracle's / LAPACK's rule on the diagonal of R
    const long long mnr = nrows < (long long)K ? nrows : (long long)K;
    const double rcond = (double)(mnr > 0 ? mnr : 1) * 2.220446049250313e-16;
    int rank = 0;
    const double r11 = mn > 0 ? fabs(M[0]) : 0.0;
    if (r11 != 0.0) for (rank = 1; rank < mn; rank++) if (!(fabs(M[rank * K + rank]) > rcond * r11)) break;
    if (lane == 0) {
        F->rank = rank;
        if (rank < K && !drop_const) F->info |= SHOPT_POLY_RANK_DEFICIENT;
    }
    if (rank == K && rc == K && !drop_const) {
        // 6a. W2 = S D Pi L^-T
        for (int i = lane; i < K * K; i += nl) Li[i] = 0.0;
        FACTOR_SYNC();
        for (int c = lane; c < K; c += nl) {              // Li = L^-1, one column per lane
            Li[c * K + c] = 1.0 / L[c * K + c];
            for (int i = c + 1; i < K; i++) {
                double s = 0.0;
                for (int k = c; k < i; k++) s -= L[i * K + k] * Li[k * K + c];
                Li[i * K + c] = s / L[i * K + i];
            }
        }
        FACTOR_SYNC();
        for (int e = lane; e < K * K; e += nl) {
            const int m = e / K, i = e % K;
            Cc[F->piv[m] * K + i] = (i >= m) ? d[F->piv[m]] * Li[i * K + m] : 0.0;
        }
        FACTOR_SYNC();
        for (int e = lane; e < K * K; e += nl) {
            const int kr = e / K, i = e % K;
            double s = 0.0;
            for (int kc = 0; kc < K; kc++) {
                const int ii = tp[kr], jj = tq[kr], p = tp[kc], q = tq[kc];
                if (ii <= p && jj <= q) s += S[kr * K + kc] * Cc[kc * K + i];
            }
            W2[kr * K + i] = s;
        }
        FACTOR_SYNC();
        return;
    }
    if (rank == 0) { FACTOR_SYNC(); return; }
    // 6b. W2 = Pm R1'(R1 R1')^-1 Qm1'
    double *Gm = Li;                                       // rank x rank, lower Cholesky in place
    for (int e = lane; e < rank * rank; e += nl) {
        const int a = e / rank, b = e % rank;
        if (b <= a) {
            double s = 0.0;
            for (int j = a; j < K; j++) s += M[a * K + j] * M[b * K + j];
            Gm[a * K + b] = s;
        }
    }
    FACTOR_SYNC();
    if (lane == 0)
        for (int a = 0; a < rank; a++)
            for (int b = 0; b <= a; b++) {
                double s = Gm[a * K + b];
                for (int t = 0; t < b; t++) s -= Gm[a * K + t] * Gm[b * K + t];
                Gm[a * K + b] = (a == b) ? sqrt(s) : s / Gm[b * K + b];
            }
    FACTOR_SYNC();
    for (int i = lane; i < rc; i += nl) {
        double cc[KMAX];
        for (int a = 0; a < rank; a++) { double s = Qm[i * K + a]; for (int t = 0; t < a; t++) s -= Gm[a * K + t] * cc[t]; cc[a] = s / Gm[a * K + a]; }
        for (int a = rank - 1; a >= 0; a--) { double s = cc[a]; for (int t = a + 1; t < rank; t++) s -= Gm[t * K + a] * cc[t]; cc[a] = s / Gm[a * K + a]; }
        for (int j = 0; j < K; j++) {
            double s = 0.0;
            for (int a = 0; a <= j && a < rank; a++) s += M[a * K + j] * cc[a];
            W2[jp[j] * K + i] = s;
        }
    }
    (void)wv;
    FACTOR_SYNC();
}

// coefficients of one right-hand side b = Vc'z through the factorisation (what solve_kernel does per pixel)
SHOPT_HD void polyfit_apply(const PolyFactor *F, const double *b, double *coef) {
    const int K = F->K, rc = F->rc_rank;
    double y[KMAX];
    for (int i = 0; i < rc; i++) {
        double s = b[F->piv[i]] * F->dsc[F->piv[i]];
        for (int k = 0; k < i; k++) s -= F->L[i * K + k] * y[k];
        y[i] = s / F->L[i * K + i];
    }
    for (int j = 0; j < K; j++) {
        double s = 0.0;
        for (int i = 0; i < rc; i++) s += F->W2[j * K + i] * y[i];
        coef[j] = s;
    }
}

}  // namespace shopt

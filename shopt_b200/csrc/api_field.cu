// api_field.cu -- s(u,v), g1(u,v), g2(u,v) across the field of view.
//
// Replaces shopt/interpolate.jl:7-87 (interpCostS / interpCostg1 / interpCostg2 and their ForwardDiff gradients) and the
// three  optimize(interpCost*, polyG_*!, rand(10), LBFGS())  calls of shopt/shopt.jl:422-466.  The cost is
//     sum_i (p1 u^3 + p2 v^3 + p3 u^2 v + p4 v^2 u + p5 u^2 + p6 v^2 + p7 u v + p8 u + p9 v + p10 - t_i)^2,
// a convex quadratic in p: whatever start LBFGS is given, its limit is the least-squares solution, which is computed
// here directly by Householder QR of the n x 10 design matrix (one factorisation shared by the three right-hand
// sides).  Host code: the work is 100 flops per star on 40 bytes per star that are already on the host.
#include <algorithm>
#include <cmath>
#include <vector>

#include "api_ctx.h"

namespace shopt {

// the reference's term order (interpolate.jl:8)
static inline void field_terms(double u, double v, double *t) {
    t[0] = u * u * u; t[1] = v * v * v; t[2] = u * u * v; t[3] = v * v * u; t[4] = u * u; t[5] = v * v; t[6] = u * v;
    t[7] = u; t[8] = v; t[9] = 1.0;
}

// min |A x - b_j| for nrhs right-hand sides; A (n x 10, column-major) and B (n x nrhs) are overwritten.
// Returns the numerical rank found on the diagonal of R (10 = full).
int field_lstsq(std::vector<double> &A, std::vector<double> &B, long long n, int nrhs, double *coef /* 10 x nrhs */) {
    constexpr int K = 10;
    int rank = 0;
    double rmax = 0.0;
    for (int k = 0; k < K && k < n; k++) {
        double *ak = A.data() + (size_t)k * n;
        double nrm = 0.0;
        for (long long i = k; i < n; i++) nrm = std::hypot(nrm, ak[i]);
        if (nrm == 0.0) continue;
        const double alpha = ak[k] > 0 ? -nrm : nrm;
        ak[k] -= alpha;                                   // v = x - alpha e1
        double vnorm2 = 0.0;
        for (long long i = k; i < n; i++) vnorm2 += ak[i] * ak[i];
        if (vnorm2 > 0.0) {
            auto reflect = [&](double *col) {
                double dot = 0.0;
                for (long long i = k; i < n; i++) dot += ak[i] * col[i];
                const double f = 2.0 * dot / vnorm2;
                for (long long i = k; i < n; i++) col[i] -= f * ak[i];
            };
            for (int j = k + 1; j < K; j++) reflect(A.data() + (size_t)j * n);
            for (int j = 0; j < nrhs; j++) reflect(B.data() + (size_t)j * n);
        }
        ak[k] = alpha;                                    // R[k][k]
        rmax = std::fmax(rmax, std::fabs(alpha));
    }
    for (int k = 0; k < K && k < n; k++)
        if (std::fabs(A[(size_t)k * n + k]) > rmax * 1e-13) rank++;
    for (int j = 0; j < nrhs; j++) {
        double *x = coef + j * K;
        for (int k = K - 1; k >= 0; k--) {
            if (k >= n || !(std::fabs(A[(size_t)k * n + k]) > rmax * 1e-13)) { x[k] = 0.0; continue; }
            double s = B[(size_t)j * n + k];
            for (int m = k + 1; m < K; m++) s -= A[(size_t)m * n + k] * x[m];
            x[k] = s / A[(size_t)k * n + k];
        }
    }
    return rank;
}


// The same least squares from its normal equations G x = b_j (G = T'T 10 x 10 row-major, b 10 x nrhs stored [j][k]):
// diagonal equilibration, Cholesky with diagonal pivoting (rank-revealing: pivots below 1e-13 of the largest end the
// factorisation and the remaining coefficients are 0).  What the sharded stage uses: G and b add over shards.
int field_solve_normal(const double *G, const double *b, int nrhs, double *coef) {
    constexpr int K = 10;
    double A[K][K], d[K];
    int perm[K];
    for (int i = 0; i < K; i++) { d[i] = G[i * K + i] > 0.0 ? 1.0 / std::sqrt(G[i * K + i]) : 0.0; perm[i] = i; }
    for (int i = 0; i < K; i++)
        for (int j = 0; j < K; j++) A[i][j] = G[i * K + j] * d[i] * d[j];
    int rank = 0;
    for (int k = 0; k < K; k++) {                          // A[perm][perm] = L L'
        int pv = k;
        for (int i = k + 1; i < K; i++) if (A[perm[i]][perm[i]] > A[perm[pv]][perm[pv]]) pv = i;
        std::swap(perm[k], perm[pv]);
        const int p = perm[k];
        if (!(A[p][p] > 1e-13)) break;
        const double l = std::sqrt(A[p][p]);
        A[p][p] = l;
        for (int i = k + 1; i < K; i++) A[perm[i]][p] /= l;
        for (int i = k + 1; i < K; i++)
            for (int j = k + 1; j <= i; j++) {
                A[perm[i]][perm[j]] -= A[perm[i]][p] * A[perm[j]][p];
                A[perm[j]][perm[i]] = A[perm[i]][perm[j]];
            }
        rank++;
    }
    for (int j = 0; j < nrhs; j++) {
        double y[K], x[K];
        for (int k = 0; k < rank; k++) {
            double s = b[j * K + perm[k]] * d[perm[k]];
            for (int m = 0; m < k; m++) s -= A[perm[k]][perm[m]] * y[m];
            y[k] = s / A[perm[k]][perm[k]];
        }
        for (int k = rank - 1; k >= 0; k--) {
            double s = y[k];
            for (int m = k + 1; m < rank; m++) s -= A[perm[m]][perm[k]] * x[m];
            x[k] = s / A[perm[k]][perm[k]];
        }
        for (int k = 0; k < K; k++) coef[j * K + k] = 0.0;
        for (int k = 0; k < rank; k++) coef[j * K + perm[k]] = x[k] * d[perm[k]];
    }
    return rank;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int shopt_field_interp(const double *u, const double *v, const double *const *values, int nvalues, int64_t n,
                       double *coef) {
    if (!u || !v || !values || !coef || nvalues < 1 || nvalues > 8 || n < 0) return SHOPT_E_BADARG;
    for (int j = 0; j < nvalues; j++) if (!values[j]) return SHOPT_E_BADARG;
    std::vector<double> A((size_t)std::max<int64_t>(n, 1) * 10), B((size_t)std::max<int64_t>(n, 1) * nvalues);
    for (int64_t i = 0; i < n; i++) {
        double t[10];
        field_terms(u[i], v[i], t);
        for (int k = 0; k < 10; k++) A[(size_t)k * n + i] = t[k];
        for (int j = 0; j < nvalues; j++) B[(size_t)j * n + i] = values[j][i];
    }
    const int rank = field_lstsq(A, B, n, nvalues, coef);
    return rank;                                          // 10 = full rank; fewer: the dependent coefficients are 0
}

int shopt_field_eval(const double *coef, const double *u, const double *v, int64_t n, double *out) {
    if (!coef || !u || !v || !out || n < 0) return SHOPT_E_BADARG;
    for (int64_t i = 0; i < n; i++) {
        double t[10], s = 0.0;
        field_terms(u[i], v[i], t);
        for (int k = 0; k < 10; k++) s += coef[k] * t[k];   // left to right like the reference's expression (shopt.jl:437)
        out[i] = s;
    }
    return SHOPT_OK;
}

}  // extern "C"

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
#include "polyfit_factor.h"

namespace shopt {

// the reference's term order (interpolate.jl:8)
static inline void field_terms(double u, double v, double *t) {
    t[0] = u * u * u; t[1] = v * v * v; t[2] = u * u * v; t[3] = v * v * u; t[4] = u * u; t[5] = v * v; t[6] = u * v;
    t[7] = u; t[8] = v; t[9] = 1.0;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

// The cubic is a degree-3 polynomial_optimizer problem in another term order: the normal equations are formed in the
// centred and scaled basis (raw sky coordinates make the raw cubic design matrix numerically singular) and solved by
// polyfit_factor.h -- the same path the stage runs on the device (api_stage.cu: field_normal_kernel).
int shopt_field_interp(const double *u, const double *v, const double *const *values, int nvalues, int64_t n,
                       double *coef) {
    if (!u || !v || !values || !coef || nvalues < 1 || nvalues > 8 || n < 0) return SHOPT_E_BADARG;
    for (int j = 0; j < nvalues; j++) if (!values[j]) return SHOPT_E_BADARG;
    constexpr int K = 10;
    double ctr[4] = {0.0, 1.0, 0.0, 1.0};
    if (n > 0) {                                          // the device's rule (select.cu: uv_centre_kernel)
        double su = 0, sv = 0, suu = 0, svv = 0;
        for (int64_t i = 0; i < n; i++) { su += u[i]; sv += v[i]; suu += u[i] * u[i]; svv += v[i] * v[i]; }
        const double u0 = su / n, v0 = sv / n, vu = suu / n - u0 * u0, vv = svv / n - v0 * v0;
        double sc_u = 1.0, sc_v = 1.0;
        if (std::isfinite(vu) && vu > 1e-12 * u0 * u0 && vu > 0.0) sc_u = std::sqrt(3.0 * vu);
        if (std::isfinite(vv) && vv > 1e-12 * v0 * v0 && vv > 0.0) sc_v = std::sqrt(3.0 * vv);
        ctr[0] = std::isfinite(u0) ? u0 : 0.0; ctr[1] = 1.0 / sc_u; ctr[2] = std::isfinite(v0) ? v0 : 0.0; ctr[3] = 1.0 / sc_v;
    }
    std::vector<double> G(K * K, 0.0), B((size_t)K * nvalues, 0.0), work(7 * K * K + 2 * K);
    for (int64_t i = 0; i < n; i++) {
        const double a = (u[i] - ctr[0]) * ctr[1], b = (v[i] - ctr[2]) * ctr[3];
        const double t[K] = {1.0, b, b * b, b * b * b, a, a * b, a * b * b, a * a, a * a * b, a * a * a};   // p outer, q inner
        for (int p = 0; p < K; p++) {
            for (int q = 0; q < K; q++) G[p * K + q] += t[p] * t[q];
            for (int j = 0; j < nvalues; j++) B[(size_t)p * nvalues + j] += t[p] * values[j][i];
        }
    }
    std::vector<PolyFactor> F(1);
    polyfit_factor(3, G.data(), ctr, n, 0, work.data(), F.data());
    const int map[K] = {9, 3, 8, 6, 7, 2, 5, 4, 1, 0};    // polynomial_optimizer's order -> interpolate.jl:8
    for (int j = 0; j < nvalues; j++) {
        double bj[KMAX], cj[KMAX];
        for (int k = 0; k < K; k++) bj[k] = B[(size_t)k * nvalues + j];
        polyfit_apply(F.data(), bj, cj);
        for (int k = 0; k < K; k++) coef[j * K + k] = cj[map[k]];
    }
    return F[0].rank;                                      // 10 = full rank; fewer: the minimum-norm solution, like A \ z
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

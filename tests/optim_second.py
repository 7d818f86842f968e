"""A SECOND, independently written restatement of  Optim.optimize(f, g!, x0, LBFGS(), Options(g_tol))  (Optim 1.7.8 +
LineSearches 7.2.0 HagerZhang + InitialStatic), in the index-into-history style of
LineSearches.jl's own source (alphas / values / slopes vectors, ia / ib indices) -- written from the published algorithm,
not from oracle/shopt_oracle.c.  TEST INFRASTRUCTURE ONLY.

Purpose: the oracle's LBFGS/HagerZhang state machine and the CUDA kernel's were written by
one author from one reading of the spec; this file is a second reading.  tests/test_optim_second.py drives it with the
ORACLE's cost and gradient evaluations (bit-identical function values), so that any difference in the optimiser logic
-- two-loop scaling rule, InitialStatic, bracketing / secant2 / update / bisect branch order, convergence-test order --
shows up as a different iterate sequence.
"""
import math

import numpy as np

EPS = float(np.finfo(np.float64).eps)


class LineSearchException(Exception):
    def __init__(self, msg, alpha):
        super().__init__(msg)
        self.alpha = alpha


class AssertionInLineSearch(Exception):
    """An @assert / error() inside LineSearches fired: propagates out of optimize (shopt.jl's catch -> zeros)."""


def _check(cond, what=""):
    if not cond:
        raise AssertionInLineSearch(what)


def eps_of(x):
    """Julia's eps(x::Float64): distance to the next larger float in magnitude."""
    x = abs(x)
    if not math.isfinite(x):
        return math.nan
    return float(np.spacing(x))


def nextfloat(x):
    return float(np.nextafter(x, math.inf))


class HagerZhang:
    delta, sigma, alphamax, rho, epsilon, gamma, linesearchmax, psi3 = 0.1, 0.9, math.inf, 5.0, 1e-6, 0.66, 50, 0.1

    def __init__(self, phidphi):
        self.phidphi = phidphi           # alpha -> (phi, dphi)
        self.trials = 0

    def _eval(self, a):
        self.trials += 1
        return self.phidphi(a)

    def satisfies_wolfe(self, c, phi_c, dphi_c, phi_0, dphi_0, phi_lim):
        wolfe1 = self.delta * dphi_0 >= (phi_c - phi_0) / c and dphi_c >= self.sigma * dphi_0
        wolfe2 = ((2 * self.delta - 1) * dphi_0 >= dphi_c >= self.sigma * dphi_0) and phi_c <= phi_lim
        return wolfe1 or wolfe2

    def __call__(self, c, phi_0, dphi_0, mayterminate=False):
        if not (math.isfinite(phi_0) and math.isfinite(dphi_0)):
            raise LineSearchException("Value and slope at step length = 0 must be finite.", 0.0)
        if dphi_0 >= EPS * abs(phi_0):
            raise LineSearchException("Search direction is not a direction of descent.", 0.0)
        elif dphi_0 >= 0:
            return 0.0, phi_0
        iterfinitemax = int(math.ceil(-math.log2(EPS)))
        self.alphas, self.values, self.slopes = [0.0], [phi_0], [dphi_0]
        alphas, values, slopes = self.alphas, self.values, self.slopes
        phi_lim = phi_0 + self.epsilon * abs(phi_0)
        alphamax = self.alphamax
        _check(c >= 0, "c >= 0")
        if c <= EPS:
            return 0.0, phi_0
        _check(math.isfinite(c) and c <= alphamax, "isfinite(c) && c <= alphamax")
        phi_c, dphi_c = self._eval(c)
        iterfinite = 1
        while not (math.isfinite(phi_c) and math.isfinite(dphi_c)) and iterfinite < iterfinitemax:
            mayterminate = False
            iterfinite += 1
            c *= self.psi3
            phi_c, dphi_c = self._eval(c)
        if not (math.isfinite(phi_c) and math.isfinite(dphi_c)):
            return 0.0, phi_0
        alphas.append(c); values.append(phi_c); slopes.append(dphi_c)
        if mayterminate and self.satisfies_wolfe(c, phi_c, dphi_c, phi_0, dphi_0, phi_lim):
            return c, phi_c
        # initial bracketing (HZ stages B0-B3); indices are 0-based here, 1-based in LineSearches.jl
        isbracketed = False
        ia, ib = 0, 1
        it = 1
        while not isbracketed and it < self.linesearchmax:
            if dphi_c >= 0.0:
                ib = len(alphas) - 1
                for i in range(ib - 1, -1, -1):
                    if values[i] <= phi_lim:
                        ia = i
                        break
                isbracketed = True
            elif values[-1] > phi_lim:
                ib = len(alphas) - 1
                ia = 0
                if (not _isapprox(c, alphas[ib])) or slopes[ib] >= 0.0:
                    raise AssertionInLineSearch("c = %r" % c)
                ia, ib = self.bisect(ia, ib, phi_lim)
                isbracketed = True
            else:
                cold, phi_cold = c, phi_c
                if nextfloat(cold) >= alphamax:
                    return cold, phi_cold
                c *= self.rho
                if c > alphamax:
                    c = alphamax
                phi_c, dphi_c = self._eval(c)
                iterfinite = 1
                while (not (math.isfinite(phi_c) and math.isfinite(dphi_c))) and c > nextfloat(cold) and iterfinite < iterfinitemax:
                    alphamax = c
                    iterfinite += 1
                    c = (cold + c) / 2
                    phi_c, dphi_c = self._eval(c)
                if not (math.isfinite(phi_c) and math.isfinite(dphi_c)):
                    return cold, phi_cold
                alphas.append(c); values.append(phi_c); slopes.append(dphi_c)
            it += 1
        while it < self.linesearchmax:
            a, b = alphas[ia], alphas[ib]
            _check(b > a, "b > a")
            if b - a <= eps_of(b):
                return a, values[ia]
            iswolfe, iA, iB = self.secant2(ia, ib, phi_lim)
            if iswolfe:
                return alphas[iA], values[iA]
            A, B = alphas[iA], alphas[iB]
            _check(B > A, "B > A")
            if B - A < self.gamma * (b - a):
                if nextfloat(values[ia]) >= values[ib] and nextfloat(values[iA]) >= values[iB]:
                    return A, values[iA]
                ia, ib = iA, iB
            else:
                c = (A + B) / 2.0
                phi_c, dphi_c = self._eval(c)
                _check(math.isfinite(phi_c) and math.isfinite(dphi_c), "finite bisection point")
                alphas.append(c); values.append(phi_c); slopes.append(dphi_c)
                ia, ib = self.update(iA, iB, len(alphas) - 1, phi_lim)
            it += 1
        raise LineSearchException("Linesearch failed to converge, reached maximum iterations.", alphas[ia])

    def secant2(self, ia, ib, phi_lim):
        alphas, values, slopes = self.alphas, self.values, self.slopes
        phi_0, dphi_0 = values[0], slopes[0]
        a, b, dphi_a, dphi_b = alphas[ia], alphas[ib], slopes[ia], slopes[ib]
        if not (dphi_a < 0.0 and dphi_b >= 0.0):
            raise AssertionInLineSearch("Search direction is not a direction of descent (secant2)")
        c = (a * dphi_b - b * dphi_a) / (dphi_b - dphi_a)
        _check(math.isfinite(c), "isfinite(c)")
        phi_c, dphi_c = self._eval(c)
        _check(math.isfinite(phi_c) and math.isfinite(dphi_c), "finite secant point")
        alphas.append(c); values.append(phi_c); slopes.append(dphi_c)
        ic = len(alphas) - 1
        if self.satisfies_wolfe(c, phi_c, dphi_c, phi_0, dphi_0, phi_lim):
            return True, ic, ic
        iA, iB = self.update(ia, ib, ic, phi_lim)
        a, b = alphas[iA], alphas[iB]
        if iB == ic:
            c = _secant(alphas, slopes, ib, iB)
        elif iA == ic:
            c = _secant(alphas, slopes, ia, iA)
        if (iA == ic or iB == ic) and a <= c <= b:
            phi_c, dphi_c = self._eval(c)
            _check(math.isfinite(phi_c) and math.isfinite(dphi_c), "finite second secant point")
            alphas.append(c); values.append(phi_c); slopes.append(dphi_c)
            ic = len(alphas) - 1
            if self.satisfies_wolfe(c, phi_c, dphi_c, phi_0, dphi_0, phi_lim):
                return True, ic, ic
            iA, iB = self.update(iA, iB, ic, phi_lim)
        return False, iA, iB

    def update(self, ia, ib, ic, phi_lim):
        alphas, values, slopes = self.alphas, self.values, self.slopes
        a, b = alphas[ia], alphas[ib]
        _check(slopes[ia] < 0.0, "slopes[ia] < 0")
        _check(values[ia] <= phi_lim, "values[ia] <= phi_lim")
        _check(slopes[ib] >= 0.0, "slopes[ib] >= 0")
        _check(b > a, "b > a")
        c, phi_c, dphi_c = alphas[ic], values[ic], slopes[ic]
        if c < a or c > b:
            return ia, ib
        if dphi_c >= 0.0:
            return ia, ic
        if phi_c <= phi_lim:
            return ic, ib
        return self.bisect(ia, ic, phi_lim)

    def bisect(self, ia, ib, phi_lim):
        alphas, values, slopes = self.alphas, self.values, self.slopes
        a, b = alphas[ia], alphas[ib]
        _check(slopes[ia] < 0.0, "bisect: slopes[ia] < 0")
        _check(values[ia] <= phi_lim, "bisect: values[ia] <= phi_lim")
        _check(slopes[ib] < 0.0, "bisect: slopes[ib] < 0")
        _check(values[ib] > phi_lim, "bisect: values[ib] > phi_lim")
        _check(b > a, "bisect: b > a")
        while b - a > eps_of(b):
            d = (a + b) / 2.0
            phi_d, gphi = self._eval(d)
            _check(math.isfinite(phi_d) and math.isfinite(gphi), "finite bisect point")
            alphas.append(d); values.append(phi_d); slopes.append(gphi)
            idd = len(alphas) - 1
            if gphi >= 0.0:
                return ia, idd
            if phi_d <= phi_lim:
                a, ia = d, idd
            else:
                b, ib = d, idd
        return ia, ib


def _secant(alphas, slopes, ia, ib):
    a, b, da, db = alphas[ia], alphas[ib], slopes[ia], slopes[ib]
    with np.errstate(all="ignore"):
        return float((np.float64(a) * db - np.float64(b) * da) / (np.float64(db) - np.float64(da)))


def _isapprox(x, y):
    """Julia's isapprox default: |x - y| <= sqrt(eps) * max(|x|, |y|)."""
    return x == y or abs(x - y) <= math.sqrt(EPS) * max(abs(x), abs(y))


CONVERGED, MAXITER, LS_FAILED, NONFINITE, THREW = 0, 1, 2, 3, -1


def dot(a, b):
    """Left-to-right sum of products, like a plain loop (numpy's BLAS dot reorders: three terms, the order matters at
    the last bit, and dphi is a sum with cancellation)."""
    t = 0.0
    for x, y in zip(a, b):
        t += float(x) * float(y)
    return t


def lbfgs(f, g, x0, g_tol=1e-8, iterations=1000, m=10, trace=None):
    """Returns (x, status, iterations, n_linesearch_trials).  f(x) -> float, g(x) -> ndarray(3)."""
    x = np.array(x0, dtype=np.float64)
    n = len(x)
    fx, gx = f(x), g(x)
    trials = 0
    if not math.isfinite(fx) or not np.all(np.isfinite(gx)):
        return x, NONFINITE, 0, 0
    if np.max(np.abs(gx)) <= g_tol:
        return x, CONVERGED, 0, 0
    dx_hist, dg_hist, rho = np.zeros((m, n)), np.zeros((m, n)), np.zeros(m)
    pseudo = 0
    counter_f_tol = 0
    it = 0
    converged = False
    status = MAXITER
    while not converged and it < iterations:
        it += 1
        # ---- update_state! ----
        pseudo += 1
        lower, upper = pseudo - m, pseudo - 1
        q = gx.copy()
        alpha_tl = np.zeros(m)
        for index in range(upper, lower - 1, -1):
            if index < 1:
                continue
            i = (index - 1) % m
            alpha_tl[i] = rho[i] * dot(dx_hist[i], q)
            q = q - alpha_tl[i] * dg_hist[i]
        if pseudo > 1:
            i = (upper - 1) % m
            scaling = dot(dx_hist[i], dg_hist[i]) / dot(dg_hist[i], dg_hist[i])
            s = scaling * q
        else:
            s = q.copy()
        for index in range(lower, upper + 1):
            if index < 1:
                continue
            i = (index - 1) % m
            beta = rho[i] * dot(dg_hist[i], s)
            s = s + dx_hist[i] * (alpha_tl[i] - beta)
        s = -s
        g_prev = gx.copy()
        # ---- perform_linesearch! ----
        dphi_0 = dot(gx, s)
        if dphi_0 >= 0.0:
            pseudo = 1
            s = -gx
            dphi_0 = dot(gx, s)
        phi_0 = fx
        f_prev, x_prev = phi_0, x.copy()
        cache = {}

        def phidphi(a, x=x, s=s):
            xn = x + a * s
            fv, gv = f(xn), g(xn)
            cache["last"] = (xn, fv, gv)
            return fv, dot(gv, s)
        ls = HagerZhang(phidphi)
        ls_ok = True
        try:
            alpha, _ = ls(1.0, phi_0, dphi_0)           # InitialStatic(alpha = 1.0); mayterminate stays false
        except LineSearchException as ex:
            alpha, ls_ok = ex.alpha, False
        except AssertionInLineSearch:
            return x, THREW, it, trials + ls.trials
        trials += ls.trials
        dx = alpha * s
        x = x + dx
        if trace is not None:
            trace.append((alpha, x.copy()))
        if not ls_ok:
            status = LS_FAILED
            break
        # ---- update_g! (value_gradient!, cached when x is the last point the line search evaluated) ----
        last = cache.get("last")
        if last is not None and np.array_equal(last[0], x):
            fx, gx = last[1], last[2]
        else:
            fx, gx = f(x), g(x)
        # ---- assess_convergence ----
        x_converged = float(np.max(np.abs(x - x_prev))) <= 0.0
        f_converged = abs(fx - f_prev) <= 0.0
        g_converged = float(np.max(np.abs(gx))) <= g_tol if np.all(np.isfinite(gx)) else False
        counter_f_tol = counter_f_tol + 1 if f_converged else 0
        converged = x_converged or g_converged or counter_f_tol > 1
        # ---- update_h! ----
        dg = gx - g_prev
        with np.errstate(divide="ignore", invalid="ignore"):
            rho_it = float(np.float64(1.0) / np.float64(dot(dx, dg)))
        if math.isinf(rho_it):
            pseudo = 0
        else:
            idx = (pseudo - 1) % m
            dx_hist[idx], dg_hist[idx], rho[idx] = dx, dg, rho_it
        if not np.all(np.isfinite(gx)):
            status = NONFINITE
            break
        if converged:
            status = CONVERGED
    return x, status, it, trials

"""ctypes loader for the CPU oracle (oracle/libshopt_oracle.so).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs, never by the product package shopt_b200/.  PARITY UNPINNED (see shopt_oracle.h).

All arrays are float64, Julia layout: a batch of N stamps of r x c pixels is a numpy array of shape
(N, c, r) C-contiguous, i.e. memory order == Julia's column-major (r, c, N); stamp pixel (u, v)
(1-based) is arr[i, v-1, u-1].
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

FIT_CONVERGED, FIT_MAXITER, FIT_LS_FAILED, FIT_NONFINITE, FIT_THREW = 0, 1, 2, 3, -1


def _cpu_stamp():
    """The library is built -march=native; a copy built on another CPU model must be rebuilt before it is loaded."""
    try:
        with open("/proc/cpuinfo") as f:
            lines = [ln.split(":", 1)[1].strip() for ln in f if ln.startswith(("model name", "flags"))][:2]
        return "|".join(lines)
    except OSError:
        return "unknown"


def build(force=False):
    so = os.path.join(_HERE, "libshopt_oracle.so")
    stamp = os.path.join(_HERE, ".build_cpu")
    src = [os.path.join(_HERE, f) for f in ("shopt_oracle.c", "shopt_oracle.h", "Makefile")]
    same_cpu = os.path.exists(stamp) and open(stamp).read() == _cpu_stamp()
    if force or not same_cpu or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
        try:
            with open(stamp, "w") as f:
                f.write(_cpu_stamp())
        except OSError:
            pass
    return so


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)


def _p(a, t=_dp):
    return None if a is None else a.ctypes.data_as(t)


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_fgaussian.restype = C.c_double
        L.orc_fgaussian.argtypes = [C.c_double] * 7
        L.orc_cost.restype = C.c_double
        L.orc_cost.argtypes = [_dp, C.c_int, C.c_int, C.c_int, _dp]
        L.orc_grad.restype = None
        L.orc_grad.argtypes = [_dp, C.c_int, C.c_int, C.c_int, _dp, _dp]
        L.orc_e_to_g.restype = None
        L.orc_e_to_g.argtypes = [C.c_double, C.c_double, _dp, _dp]
        L.orc_lbfgs_fit.restype = C.c_int
        L.orc_lbfgs_fit.argtypes = [_dp, C.c_int, C.c_int, C.c_int, _dp, C.c_double, C.c_int, _dp, _dp,
                                    C.POINTER(C.c_int)]
        L.orc_fit_analytic_batch.restype = None
        L.orc_fit_analytic_batch.argtypes = [_dp, C.c_int, C.c_int, C.c_int64, C.c_int, _dp, C.c_double,
                                             C.c_int, _dp, _dp, _dp, _ip, _ip, C.c_int]
        L.orc_s_out_of_bounds.restype = C.c_int64
        L.orc_s_out_of_bounds.argtypes = [_dp, C.c_int64, C.c_double, C.c_double, _lp]
        L.orc_nan_to_zero.restype = None
        L.orc_nan_to_zero.argtypes = [_dp, C.c_int64, _dp]
        L.orc_pca_image.restype = C.c_int
        L.orc_pca_image.argtypes = [_dp, C.c_int, C.c_int, C.c_int, C.c_double, _dp, C.POINTER(C.c_int)]
        L.orc_lanczos_kernel2d.restype = None
        L.orc_lanczos_kernel2d.argtypes = [C.c_int, _dp]
        L.orc_smooth.restype = None
        L.orc_smooth.argtypes = [_dp, C.c_int, C.c_int, C.c_int, _dp]
        L.orc_pixel_grid_pca_batch.restype = None
        L.orc_pixel_grid_pca_batch.argtypes = [_dp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_double, C.c_int,
                                               C.c_int, _dp, _ip, _ip, C.c_int]
        L.orc_num_terms.restype = C.c_int
        L.orc_num_terms.argtypes = [C.c_int]
        L.orc_objective_function.restype = C.c_double
        L.orc_objective_function.argtypes = [_dp, C.c_double, C.c_double, C.c_int]
        L.orc_polynomial_optimizer.restype = None
        L.orc_polynomial_optimizer.argtypes = [C.c_int, _dp, _dp, _dp, C.c_int64, C.c_int, _dp]
        L.orc_median.restype = C.c_double
        L.orc_median.argtypes = [_dp, C.c_int64]
        L.orc_worst_10_percent.restype = C.c_int64
        L.orc_worst_10_percent.argtypes = [_dp, C.c_int64, C.c_int, _lp]
        L.orc_polyfit.restype = None
        L.orc_polyfit.argtypes = [_dp, C.c_int, C.c_int, C.c_int64, _dp, _dp, _lp, C.c_int64, C.c_int, C.c_int,
                                  C.c_int, C.c_int, _dp, _dp, _lp, _lp, _lp, C.c_int]
        L.orc_psf_at.restype = None
        L.orc_psf_at.argtypes = [_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _dp]
        L.orc_max_threads.restype = C.c_int
        L.orc_fkolmogorov.restype = C.c_double
        L.orc_fkolmogorov.argtypes = [C.c_double] * 7
        L.orc_set_profile.restype = None
        L.orc_set_profile.argtypes = [C.c_int]
        L.orc_get_profile.restype = C.c_int
        L.orc_get_profile.argtypes = []
        L.orc_set_reference_shaped.restype = None
        L.orc_set_reference_shaped.argtypes = [C.c_int]
        _LIB = L
    return _LIB


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def max_threads():
    return lib().orc_max_threads()


def set_reference_shaped(on):
    lib().orc_set_reference_shaped(int(bool(on)))


def set_profile(profile):
    """0: fGaussian (default), 1: fkolmogorov -- the profile orc_cost / orc_grad / lbfgs_fit evaluate."""
    lib().orc_set_profile(int(profile))


def fkolmogorov(u, v, g1, g2, s, uc, vc):
    return lib().orc_fkolmogorov(u, v, g1, g2, s, uc, vc)


def fgaussian(u, v, g1, g2, s, uc, vc):
    return lib().orc_fgaussian(u, v, g1, g2, s, uc, vc)


def e_to_g(e1, e2):
    a, b = C.c_double(), C.c_double()
    lib().orc_e_to_g(e1, e2, C.byref(a), C.byref(b))
    return a.value, b.value


def cost(stamp, n_a, x):
    stamp = _f64(stamp); x = _f64(x)
    c, r = stamp.shape
    return lib().orc_cost(_p(stamp), r, c, n_a, _p(x))


def grad(stamp, n_a, x):
    stamp = _f64(stamp); x = _f64(x)
    c, r = stamp.shape
    g = np.zeros(3)
    lib().orc_grad(_p(stamp), r, c, n_a, _p(x), _p(g))
    return g


def lbfgs_fit(stamp, n_a, x0, g_tol=1e-6, max_iter=1000):
    stamp = _f64(stamp); x0 = _f64(x0)
    c, r = stamp.shape
    ox, op = np.zeros(3), np.zeros(3)
    cnt = (C.c_int * 3)()
    st = lib().orc_lbfgs_fit(_p(stamp), r, c, n_a, _p(x0), g_tol, max_iter, _p(ox), _p(op), cnt)
    return dict(status=st, x=ox, p=op, iterations=cnt[0], evals=cnt[1], trials=cnt[2])


def fit_analytic_batch(stamps, n_a, x0, g_tol=1e-6, max_iter=1000, nthreads=1):
    stamps = _f64(stamps); x0 = _f64(x0)
    N, c, r = stamps.shape
    s, g1, g2 = np.zeros(N), np.zeros(N), np.zeros(N)
    status, iters = np.zeros(N, np.int32), np.zeros(N, np.int32)
    lib().orc_fit_analytic_batch(_p(stamps), r, c, N, n_a, _p(x0), g_tol, max_iter, _p(s), _p(g1), _p(g2),
                                 _p(status, _ip), _p(iters, _ip), nthreads)
    return s, g1, g2, status, iters


def s_out_of_bounds(s, s_lo, s_hi):
    s = _f64(s)
    idx = np.zeros(len(s), np.int64)
    n = lib().orc_s_out_of_bounds(_p(s), len(s), s_lo, s_hi, _p(idx, _lp))
    return idx[:n].copy()


def nan_to_zero(a):
    a = _f64(a); out = np.empty_like(a)
    lib().orc_nan_to_zero(_p(a), a.size, _p(out))
    return out


def pca_image(img, ncomp, pratio=0.99):
    img = _f64(img)
    c, r = img.shape
    out = np.zeros_like(img); k = C.c_int()
    st = lib().orc_pca_image(_p(img), r, c, ncomp, pratio, _p(out), C.byref(k))
    return out, k.value, st


def lanczos_kernel2d(a):
    k = np.zeros((2 * a + 1, 2 * a + 1))
    lib().orc_lanczos_kernel2d(a, _p(k))
    return k


def smooth(img, a):
    img = _f64(img)
    c, r = img.shape
    out = np.zeros_like(img)
    lib().orc_smooth(_p(img), r, c, a, _p(out))
    return out


def pixel_grid_pca_batch(stamps, ncomp=50, pratio=0.99, lanczos_a=3, unity_sum=False, nthreads=1):
    stamps = _f64(stamps)
    N, c, r = stamps.shape
    pgf = np.zeros_like(stamps)
    k = np.zeros(N, np.int32); st = np.zeros(N, np.int32)
    lib().orc_pixel_grid_pca_batch(_p(stamps), r, c, N, ncomp, pratio, lanczos_a, int(unity_sum), _p(pgf),
                                   _p(k, _ip), _p(st, _ip), nthreads)
    return pgf, k, st


def num_terms(degree):
    return lib().orc_num_terms(degree)


def objective_function(coef, x, y, degree):
    coef = _f64(coef)
    return lib().orc_objective_function(_p(coef), x, y, degree)


def polynomial_optimizer(degree, x, y, z, median_constraint=False):
    x = _f64(x); y = _f64(y); z = np.array(z, dtype=np.float64)
    coef = np.zeros(num_terms(degree))
    lib().orc_polynomial_optimizer(degree, _p(x), _p(y), _p(z), len(z), int(median_constraint), _p(coef))
    return coef


def median(z):
    z = _f64(z)
    return lib().orc_median(_p(z), len(z))


def worst_10_percent(err, sigma_clip=False):
    err = _f64(err)
    idx = np.zeros(max(len(err), 1), np.int64)
    n = lib().orc_worst_10_percent(_p(err), len(err), int(sigma_clip), _p(idx, _lp))
    return idx[:n].copy()


def polyfit(pgf, u, v, train_idx, degree, iterations=3, median_constraint=False, sigma_clip=False, nthreads=1):
    pgf = _f64(pgf); u = _f64(u); v = _f64(v)
    train_idx = np.ascontiguousarray(train_idx, dtype=np.int64)
    N, c, r = pgf.shape
    K = num_terms(degree)
    nt = len(train_idx)
    poly = np.zeros((K, c, r))
    mse = np.zeros(max(nt, 1)); kept = np.zeros(max(nt, 1), np.int64)
    n_kept = C.c_int64(); n_per_round = np.zeros(max(iterations, 1), np.int64)
    lib().orc_polyfit(_p(pgf), r, c, N, _p(u), _p(v), _p(train_idx, _lp), nt, degree, iterations,
                      int(median_constraint), int(sigma_clip), _p(poly), _p(mse), _p(kept, _lp),
                      C.byref(n_kept), _p(n_per_round, _lp), nthreads)
    nk = n_kept.value
    n_last = int(n_per_round[iterations - 1]) if iterations > 0 else 0
    return dict(poly=poly, mse=mse[:n_last].copy(), kept=kept[:nk].copy(), n_per_round=n_per_round.copy())


def psf_at(poly, degree, u, v):
    poly = _f64(poly)
    K, c, r = poly.shape
    star = np.zeros((c, r))
    lib().orc_psf_at(_p(poly), r, c, degree, u, v, _p(star))
    return star

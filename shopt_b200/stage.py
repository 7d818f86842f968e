"""Host-side mirror of the reference's per-star fitting stage (the loops of shopt/shopt.jl), one batched call
per loop, on top of libshopt_b200.so.

Two flavours of every call:
  * numpy arrays in host memory  -> the library's host entry points (H2D, kernels, D2H inside the call);
  * torch CUDA tensors           -> the `_dev` entry points on the current torch stream (no copies).
torch is used for device memory and streams only.  Arrays follow Julia's memory order: a batch of N stamps
of r x c pixels has shape (N, c, r), C-contiguous, so that pixel (u, v) (1-based) is arr[i, v-1, u-1].
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, dp, ip, lp


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a, t=dp):
    return None if a is None else a.ctypes.data_as(t)


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev_check(t, dtype):
    import torch
    assert t.is_cuda and t.is_contiguous() and t.dtype == dtype, "need a contiguous CUDA tensor of " + str(dtype)
    assert t.device.index == torch.cuda.current_device(), "tensor is not on the current device"
    return C.c_void_p(t.data_ptr())


# ------------------------------------------------------------------------------------------------
# (1) analytic profile fit: shopt.jl:130-164 (model stars) / :357-397 (learned stars)
# ------------------------------------------------------------------------------------------------
def fit_analytic_batch(stamps, n_a, x0, g_tol=1e-6, max_iter=1000, profile=0):
    """Optim.optimize(cost, g!, x0, LBFGS(), Options(g_tol)) for every star of the batch.
    profile: 0 = fGaussian (what the reference's cost calls), 1 = fkolmogorov (radialProfiles.jl:11-24, closed form).

    stamps (N, c, r); x0 (N, 3) start points [sigma, e1, e2] (the reference draws rand(3), shopt.jl:134).
    Returns s, g1, g2, status, iters.  Failed stars (status == FIT_THREW) carry s = g1 = g2 = 0 like the
    reference's catch (shopt.jl:155-162).
    """
    L = _lib.lib()
    if _is_torch(stamps):
        import torch
        N, c, r = stamps.shape
        dev = stamps.device
        s = torch.empty(N, dtype=torch.float64, device=dev)
        g1 = torch.empty_like(s)
        g2 = torch.empty_like(s)
        status = torch.empty(N, dtype=torch.int32, device=dev)
        iters = torch.empty_like(status)
        check(L.shopt_fit_profile_batch_dev(profile, _dev_check(stamps, torch.float64), r, c, N, n_a,
                                             _dev_check(x0, torch.float64), g_tol, max_iter,
                                             _dev_check(s, torch.float64), _dev_check(g1, torch.float64),
                                             _dev_check(g2, torch.float64), _dev_check(status, torch.int32),
                                             _dev_check(iters, torch.int32), _stream()))
        return s, g1, g2, status, iters
    stamps = _f64(stamps)
    x0 = _f64(x0)
    N, c, r = stamps.shape
    assert x0.shape == (N, 3)
    s, g1, g2 = np.zeros(N), np.zeros(N), np.zeros(N)
    status, iters = np.zeros(N, np.int32), np.zeros(N, np.int32)
    fit = L.shopt_fit_profile_batch if profile else L.shopt_fit_analytic_batch      # both spellings stay exercised
    args = (_ptr(stamps), r, c, N, n_a, _ptr(x0), g_tol, max_iter, _ptr(s), _ptr(g1), _ptr(g2), _ptr(status, ip),
            _ptr(iters, ip))
    check(fit(profile, *args) if profile else fit(*args))
    return s, g1, g2, status, iters


def analytic_cost_grad_batch(stamps, n_a, x, profile=0):
    """cost(x_i) and g!(., x_i) of analyticLBFGS.jl:8-46, 89-94 for every star i; host arrays. Returns f (N), g (N,3)."""
    stamps = _f64(stamps)
    x = _f64(x)
    N, c, r = stamps.shape
    assert x.shape == (N, 3)
    f, g = np.zeros(N), np.zeros((N, 3))
    if profile:
        check(_lib.lib().shopt_profile_cost_grad_batch(profile, _ptr(stamps), r, c, N, n_a, _ptr(x), _ptr(f), _ptr(g)))
    else:
        check(_lib.lib().shopt_analytic_cost_grad_batch(_ptr(stamps), r, c, N, n_a, _ptr(x), _ptr(f), _ptr(g)))
    return f, g


def s_out_of_bounds(s, s_lo=0.075, s_hi=1.0):
    """s-blacklist of shopt.jl:166-171 / failed list of :384-386; 0-based indices, ascending."""
    s = _f64(s)
    idx = np.zeros(max(len(s), 1), np.int64)
    n = _lib.lib().shopt_s_out_of_bounds(_ptr(s), len(s), s_lo, s_hi, _ptr(idx, lp))
    if n < 0:
        check(int(n))
    return idx[:n].copy()


def last_timing():
    """(h2d_ms, kernel_ms, d2h_ms) of the last host-buffer call, from CUDA events on the library's stream."""
    ms = (C.c_double * 3)()
    check(_lib.lib().shopt_last_timing(ms))
    return tuple(ms)


# ------------------------------------------------------------------------------------------------
# (3) per-pixel polynomial interpolation with outlier-star rejection: shopt.jl:492-577
# ------------------------------------------------------------------------------------------------
def num_terms(degree):
    return (degree + 1) * (degree + 2) // 2


def polyfit(pgf, u, v, train_idx, degree, iterations=3, median_constraint=False, sigma_clip=False):
    """`iterations` rounds of polynomial_optimizer per pixel (interpolate.jl:105-147), per-star MSE and
    worst_10_percent rejection (shopt.jl:513-575).  pgf (N, c, r) host array; train_idx: 0-based star indices
    (the reference's randperm split, shopt.jl:472-490, is an input).
    Returns dict(poly (K, c, r), mse, kept, n_per_round) like the oracle's driver."""
    pgf = _f64(pgf); u = _f64(u); v = _f64(v)
    train_idx = np.ascontiguousarray(train_idx, dtype=np.int64)
    N, c, r = pgf.shape
    K = num_terms(degree)
    nt = len(train_idx)
    poly = np.zeros((K, c, r))
    mse = np.zeros(max(nt, 1)); kept = np.zeros(max(nt, 1), np.int64)
    n_kept = C.c_int64(); npr = np.zeros(max(iterations, 1), np.int64)
    check(_lib.lib().shopt_polyfit(_ptr(pgf), r, c, N, _ptr(u), _ptr(v), _ptr(train_idx, lp), nt, degree, iterations,
                                   int(median_constraint), int(sigma_clip), _ptr(poly), _ptr(mse), _ptr(kept, lp),
                                   C.byref(n_kept), _ptr(npr, lp)))
    n_last = int(npr[iterations - 1]) if iterations > 0 else 0
    return dict(poly=poly, mse=mse[:n_last].copy(), kept=kept[:n_kept.value].copy(), n_per_round=npr[:iterations].copy())


def worst_10_percent(err, sigma_clip=False):
    """shopt.jl:513-532: positions of the worst ceil(10 %) errors (stable descending) or the 3-sigma variant."""
    err = _f64(err)
    idx = np.zeros(max(len(err), 1), np.int64)
    n = _lib.lib().shopt_worst_10_percent(_ptr(err), len(err), int(sigma_clip), _ptr(idx, lp))
    if n < 0:
        check(int(n))
    return idx[:n].copy()


def psf_at_batch(poly, degree, u, v):
    """reader.jl:40-68 p(u, v, polMatrix, degree) for M positions: (M, c, r), each stamp divided by its sum."""
    poly = _f64(poly); u = _f64(np.atleast_1d(u)); v = _f64(np.atleast_1d(v))
    K, c, r = poly.shape
    assert K == num_terms(degree) and u.shape == v.shape
    out = np.zeros((len(u), c, r))
    check(_lib.lib().shopt_psf_at_batch(_ptr(poly), r, c, degree, _ptr(u), _ptr(v), len(u), _ptr(out)))
    return out


# ------------------------------------------------------------------------------------------------
# (2) pixel-grid fit: shopt.jl:281-314 (mode "PCA") and :316-348 (mode "smooth")
# ------------------------------------------------------------------------------------------------
def pca_batch(stamps, ncomp=50, pratio=0.99, lanczos_a=3, unity_sum=False):
    """nanToZero -> pca_image(., ncomp) -> smooth(., lanczos_a) (or ./sum when unity_sum) per star.
    stamps (N, c, r) host array or CUDA tensor.  Returns pgf (N, c, r), k (components kept), status (PCA_*);
    failed stars carry zeros like the reference's catch (shopt.jl:305-309)."""
    L = _lib.lib()
    if _is_torch(stamps):
        import torch
        N, c, r = stamps.shape
        pgf = torch.empty_like(stamps)
        k = torch.empty(N, dtype=torch.int32, device=stamps.device)
        status = torch.empty_like(k)
        check(L.shopt_pca_batch_dev(_dev_check(stamps, torch.float64), r, c, N, ncomp, pratio, lanczos_a, int(unity_sum),
                                    _dev_check(pgf, torch.float64), _dev_check(k, torch.int32),
                                    _dev_check(status, torch.int32), _stream()))
        return pgf, k, status
    stamps = _f64(stamps)
    N, c, r = stamps.shape
    pgf = np.zeros_like(stamps)
    k, status = np.zeros(N, np.int32), np.zeros(N, np.int32)
    check(L.shopt_pca_batch(_ptr(stamps), r, c, N, ncomp, pratio, lanczos_a, int(unity_sum), _ptr(pgf), _ptr(k, ip),
                            _ptr(status, ip)))
    return pgf, k, status


def smooth_batch(stamps, lanczos_a=3, nan_to_zero=False):
    """smooth(data, a) of dataPreprocessing.jl:112-139 per star (mode "smooth": nan_to_zero=True)."""
    stamps = _f64(stamps)
    N, c, r = stamps.shape
    out = np.zeros_like(stamps)
    check(_lib.lib().shopt_smooth_batch(_ptr(stamps), r, c, N, lanczos_a, int(nan_to_zero), _ptr(out)))
    return out


def smooth_mode_batch(stamps, lanczos_a=3, unity_sum=False):
    """Loop body of mode "smooth" (shopt.jl:316-348): smooth(nanToZero(star), lanczos) or, with unity_sum,
    nanToZero(star) ./ sum (the reference's 1-argument smooth() raises, so nothing is smoothed)."""
    stamps = _f64(stamps)
    N, c, r = stamps.shape
    out = np.zeros_like(stamps)
    check(_lib.lib().shopt_smooth_mode_batch(_ptr(stamps), r, c, N, lanczos_a, int(unity_sum), _ptr(out)))
    return out


def chisq_batch(stamps, weights, x0=None, fill=0.0, unity_sum=False):
    """Mode "chisq" (chisq.jl:1-11, shopt.jl:194-224): the minimiser of sum(isfinite(t) ? t : 0), t = (p - star)^2 / w.
    x0 (N, c, r): start values (the reference's rand(r*c)), kept by pixels without a finite convex term."""
    stamps, weights = _f64(stamps), _f64(weights)
    N, c, r = stamps.shape
    out = np.zeros_like(stamps)
    x0 = None if x0 is None else _f64(x0)
    check(_lib.lib().shopt_chisq_batch(_ptr(stamps), _ptr(weights), _ptr(x0) if x0 is not None else None, float(fill),
                                       r, c, N, int(unity_sum), _ptr(out)))
    return out


def field_interp(u, v, *values):
    """s(u,v), g1(u,v), g2(u,v): cubic least squares in the reference's coefficient order
    [u^3, v^3, u^2 v, v^2 u, u^2, v^2, u v, u, v, 1] (interpolate.jl:7-87, shopt.jl:422-466).
    Returns coef (len(values), 10) and the numerical rank."""
    import ctypes as C
    u, v = _f64(u), _f64(v)
    vals = [_f64(x) for x in values]
    arr = (dp * len(vals))(*[_ptr(x) for x in vals])
    coef = np.zeros((len(vals), 10))
    rank = _lib.lib().shopt_field_interp(_ptr(u), _ptr(v), arr, len(vals), len(u), _ptr(coef))
    if rank < 0:
        check(rank)
    return coef, rank


def field_eval(coef, u, v):
    coef, u, v = _f64(coef), _f64(u), _f64(v)
    out = np.zeros(len(u))
    check(_lib.lib().shopt_field_eval(_ptr(coef), _ptr(u), _ptr(v), len(u), _ptr(out)))
    return out


# ------------------------------------------------------------------------------------------------
# (0) ingest: cataloging() of dataPreprocessing.jl:342-397 without the FITS decoding
# ------------------------------------------------------------------------------------------------
def ingest_batch(vignets, unity_sum=False, new_dim=None, lanczos_a=3):
    """vignets (N, c0, r0) float32 or float64 raw VIGNET cube -> stamps (N, nd, nd) float64: nanMask, optional
    unity-sum, crop / bilinear resample to new_dim with renormalisation + smooth (see include/shopt_b200.h)."""
    v = np.ascontiguousarray(vignets)
    if v.dtype != np.float32:
        v = np.ascontiguousarray(v, dtype=np.float64)
    N, c0, r0 = v.shape
    nd = r0 if new_dim is None else int(new_dim)
    out = np.zeros((N, nd, nd) if nd != r0 else (N, c0, r0))
    check(_lib.lib().shopt_ingest_batch(v.ctypes.data, int(v.dtype == np.float32), r0, c0, N, int(unity_sum), nd,
                                        lanczos_a, _ptr(out)))
    return out


def snr_cut(snr, k):
    """outliers_filter of outliers.jl:16-34: (0-based indices with snr < quantile(snr, k), the quantile)."""
    import ctypes as C
    snr = np.ascontiguousarray(snr, dtype=np.float32)
    idx = np.zeros(max(len(snr), 1), np.int64)
    q1 = C.c_double()
    n = _lib.lib().shopt_snr_cut(snr.ctypes.data_as(C.POINTER(C.c_float)), len(snr), float(k), _ptr(idx, lp), C.byref(q1))
    if n < 0:
        check(int(n))
    return idx[:n].copy(), q1.value


# ------------------------------------------------------------------------------------------------
# summary.shopt (dataOutprocessing.jl:52-84, reader.jl:13-38)
# ------------------------------------------------------------------------------------------------
def write_summary(path, poly, degree, field):
    """poly (K, c, r); field (3, 10) = s, g1, g2 cubic coefficients.  Writes the truncated summary.shopt."""
    poly, field = _f64(poly), _f64(field)
    K, c, r = poly.shape
    check(_lib.lib().shopt_summary_write(str(path).encode(), _ptr(poly), r, c, degree, _ptr(field[0]), _ptr(field[1]),
                                         _ptr(field[2])))


def write_summary_full(path, poly, degree, field, u, v, model, learned, stars, pgf, train_u, train_v, valid_stars, valid_u,
                       valid_v, flag_snr=(), flag_blacklist=(), flag_failed=()):
    """The untruncated summary.shopt (dataOutprocessing.jl:86-168).  poly (K, c, r); field (3, 10); model / learned (3, n)
    = s, g1, g2 per star; stars / pgf (n, c, r); valid_stars (n_valid, c, r); flags: 1-based star numbers."""
    import ctypes as C
    keep = []

    def arr(a):
        a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
        keep.append(a)
        return _ptr(a) if a.size else None
    poly, field, model, learned = _f64(poly), _f64(field), _f64(model), _f64(learned)
    stars, pgf, valid_stars = _f64(stars), _f64(pgf), _f64(valid_stars)
    K, c, r = poly.shape
    f = _lib.SummaryFull()
    f.poly, f.r, f.c, f.degree = _ptr(poly), r, c, degree
    f.sC, f.g1C, f.g2C = _ptr(field[0]), _ptr(field[1]), _ptr(field[2])
    f.n_stars = stars.shape[0]
    f.u, f.v = arr(u), arr(v)
    f.s_model, f.g1_model, f.g2_model = arr(model[0]), arr(model[1]), arr(model[2])
    f.s_data, f.g1_data, f.g2_data = arr(learned[0]), arr(learned[1]), arr(learned[2])
    f.stars, f.pgf = arr(stars), arr(pgf)
    f.n_train, f.train_u, f.train_v = len(train_u), arr(train_u), arr(train_v)
    f.n_valid, f.valid_u, f.valid_v, f.valid_stars = valid_stars.shape[0], arr(valid_u), arr(valid_v), arr(valid_stars)
    f.n_flag_snr, f.n_flag_blacklist, f.n_flag_failed = len(flag_snr), len(flag_blacklist), len(flag_failed)
    f.flag_snr, f.flag_blacklist, f.flag_failed = arr(flag_snr), arr(flag_blacklist), arr(flag_failed)
    check(_lib.lib().shopt_summary_write_full(str(path).encode(), C.byref(f)))


def read_summary(path):
    """read_shopt of reader.jl:13-38: (poly (K, c, r), degree, field (3, 10))."""
    import ctypes as C
    r, c, deg = C.c_int(), C.c_int(), C.c_int()
    L = _lib.lib()
    check(L.shopt_summary_read(str(path).encode(), None, 0, C.byref(r), C.byref(c), C.byref(deg), None, None, None))
    K = num_terms(deg.value)
    poly = np.zeros((K, c.value, r.value))
    field = np.zeros((3, 10))
    check(L.shopt_summary_read(str(path).encode(), _ptr(poly), poly.size, C.byref(r), C.byref(c), C.byref(deg),
                               _ptr(field[0]), _ptr(field[1]), _ptr(field[2])))
    return poly, deg.value, field

"""CPU restatement of the whole per-star fitting stage, composed from the oracle's pieces in the order of
shopt/shopt.jl (:130-184 model fit + s-blacklist, :281-314 PCA pixel-grid fit, :357-417 learned fit + failed list,
:472-490 training split, :534-577 polyfit rounds).  TEST INFRASTRUCTURE ONLY (see shopt_oracle.h); PARITY UNPINNED.

Stars keep their original indices here (flags instead of splices), like the product's shopt_stage_run; the training
set is the first round(training_ratio * n_survivors) survivors of `perm`.
"""
import numpy as np

from . import shopt_oracle as orc

FLAG_BLACKLISTED, FLAG_PCA_FAILED, FLAG_LEARNED_OOB, FLAG_TRAINING, FLAG_REJECTED = 1, 2, 4, 8, 16


def nan_mask(a):
    """masks.jl:5-17"""
    out = np.array(a, dtype=np.float64)
    out[out < -1000] = np.nan
    return out


def get_middle_nxn(a, n):
    """dataPreprocessing.jl:207-216 on an array stored (c, r): both starts are div(size - n, 2) (0-based)."""
    cs, rs = (a.shape[0] - n) // 2, (a.shape[1] - n) // 2
    return a[cs:cs + n, rs:rs + n]


def oversample_image(img, nd):
    """dataPreprocessing.jl:141-162 (bilinear); img stored (c, r) = img[col, row], so image[y, x] of the reference
    (y = row index, first) is img[x-1, y-1]."""
    size1, size2 = img.shape[1], img.shape[0]
    scale = 1 / (nd / size1)
    out = np.zeros((nd, nd))
    for y in range(1, nd + 1):
        for x in range(1, nd + 1):
            sx, sy = (x - 0.5) * scale + 0.5, (y - 0.5) * scale + 0.5
            x0 = max(1, int(np.floor(sx))); x1 = min(size1, x0 + 1)
            y0 = max(1, int(np.floor(sy))); y1 = min(size2, y0 + 1)
            dx, dy = sx - x0, sy - y0
            I = lambda yy, xx: img[xx - 1, yy - 1]
            out[x - 1, y - 1] = ((1 - dx) * (1 - dy) * I(y0, x0) + dx * (1 - dy) * I(y0, x1) + (1 - dx) * dy * I(y1, x0)
                                 + dx * dy * I(y1, x1))
    return out


def ingest(vignets, unity_sum=False, new_dim=None, lanczos=3):
    """The per-star part of cataloging() (dataPreprocessing.jl:342-397): nanMask, optional unity sum, and for
    new_dim != size the crop / oversample followed by nm(.)/sum(nz(nm(.))) and smooth(., lanczos)."""
    out = []
    for img in vignets:
        a = nan_mask(img)
        if unity_sum:
            a = a / orc.nan_to_zero(a).sum()
        r0 = a.shape[1]
        if new_dim is not None and new_dim != r0:
            b = get_middle_nxn(a, new_dim) if new_dim < r0 else oversample_image(a, new_dim)
            b = nan_mask(b)
            b = b / orc.nan_to_zero(b).sum()
            a = orc.smooth(np.ascontiguousarray(b), lanczos) if lanczos > 0 else b
        out.append(a)
    return np.stack(out)


def snr_cut(snr, k):
    """outliers.jl:16-34: indices with snr < quantile(snr, k) (Statistics' type-7 quantile of a Float32 vector)."""
    v = np.sort(np.asarray(snr, dtype=np.float32))
    n = len(v)
    aleph = n * k + (1.0 - k)
    j = int(min(max(np.trunc(aleph), 1), max(n - 1, 1)))
    gamma = min(max(aleph - j, 0.0), 1.0)
    a, b = v[j - 1], v[j] if n > 1 else v[0]
    q1 = float(a) + gamma * float(np.float32(b - a))
    return np.nonzero(np.asarray(snr, dtype=np.float32).astype(np.float64) < q1)[0], q1


def field_interp(u, v, *values):
    """interpolate.jl:7-87: the minimiser of sum (I_p(u,v) - t)^2 over the 10 cubic coefficients, term order
    [u^3, v^3, u^2 v, v^2 u, u^2, v^2, u v, u, v, 1] (numpy lstsq: SVD-based, independent of the product's QR)."""
    A = np.stack([u ** 3, v ** 3, u * u * v, v * v * u, u * u, v * v, u * v, u, v, np.ones_like(u)], axis=1)
    return np.stack([np.linalg.lstsq(A, t, rcond=None)[0] for t in values])


def pixel_grid_elementwise(mode, stamps, lanczos, unity_sum, weights=None, x0_pixels=None, fill=0.0):
    """mode 1 = "smooth" (shopt.jl:316-348), mode 2 = "chisq" (shopt.jl:194-224 with chisq.jl:1-11; the minimiser of
    the separable objective: star where the term is a finite convex quadratic, the start value elsewhere)."""
    out = np.empty_like(stamps)
    for i in range(stamps.shape[0]):
        if mode == 1:
            data = orc.nan_to_zero(stamps[i])
            out[i] = data / data.sum() if unity_sum else orc.smooth(data, lanczos)   # 1-arg smooth() raises -> :330
        else:
            w, d = weights[i], stamps[i]
            live = np.isfinite(d) & np.isfinite(w) & (w > 0)
            start = x0_pixels[i] if x0_pixels is not None else np.full_like(d, fill)
            m = np.where(live, d, start)
            out[i] = m / m.sum() if unity_sum else m
    return out


def run_stage(p, stamps, u, v, x0_model, x0_learned, perm, nthreads=1, fit_only=False, weights=None, x0_pixels=None,
              nthreads_poly=None):
    """p: any object with the shopt_params fields.  Returns the same dict layout as shopt_b200.pipeline.run_stage.
    nthreads: threads of the per-star loops (the reference runs them serially); nthreads_poly: threads of the per-pixel
    polyfit (the reference's only @threads loop, shopt.jl:543); default: nthreads."""
    N = stamps.shape[0]
    nthreads_poly = nthreads if nthreads_poly is None else nthreads_poly
    orc.set_profile(getattr(p, "profile", 0))     # radial profile of both analytic fits (0: the reference's fGaussian)
    s, g1, g2, mst, mit = orc.fit_analytic_batch(stamps, p.n_a, x0_model, p.g_tol_model, p.max_iter, nthreads)
    model = np.stack([s, g1, g2])
    if fit_only:
        return dict(model=model, model_status=mst, model_iters=mit)
    flags = np.zeros(N, np.int32)
    flags[(s < p.s_lo) | (s > p.s_hi)] |= FLAG_BLACKLISTED                         # shopt.jl:166-171
    mode = getattr(p, "mode", 0)
    if mode == 0:                                                                    # shopt.jl:281-314
        pgf, k, pst = orc.pixel_grid_pca_batch(stamps, p.pca_terms, p.pratio, p.lanczos, bool(p.unity_sum), nthreads)
    else:
        pgf = pixel_grid_elementwise(mode, stamps, p.lanczos, bool(p.unity_sum), weights, x0_pixels,
                                     getattr(p, "chisq_fill", 0.0))
        k, pst = np.zeros(N, np.int32), np.zeros(N, np.int32)
    flags[pst == 1] |= FLAG_PCA_FAILED                                               # shopt.jl:305-309
    sl, g1l, g2l, lst, lit = orc.fit_analytic_batch(pgf, p.n_a, x0_learned, p.g_tol_learned, p.max_iter, nthreads)
    # shopt.jl:384-386 sits INSIDE the try: a learned fit that threw lands in the catch (:387-393: s = g1 = g2 = 0) and
    # is never pushed to failedStars
    flags[((sl < p.s_lo) | (sl > p.s_hi)) & (lst != orc.FIT_THREW)] |= FLAG_LEARNED_OOB
    surv = flags == 0
    field = field_interp(u[surv], v[surv], sl[surv], g1l[surv], g2l[surv])        # shopt.jl:422-466
    n_train = int(np.rint(p.training_ratio * surv.sum()))                            # round(Int, .): half to even
    train = np.array([g for g in perm if surv[g]][:n_train], dtype=np.int64)
    flags[train] |= FLAG_TRAINING
    res = orc.polyfit(pgf, u, v, train, p.degree, p.iterations, bool(p.median_constraint), bool(p.sigma_clip), nthreads_poly)
    rejected = np.setdiff1d(train, res["kept"])
    flags[rejected] |= FLAG_REJECTED
    black = (flags & FLAG_BLACKLISTED) != 0
    failed = ~black & ((flags & (FLAG_PCA_FAILED | FLAG_LEARNED_OOB)) != 0)
    return dict(model=model, learned=np.stack([sl, g1l, g2l]), pgf=pgf, flags=flags, pca_k=k, poly=res["poly"],
                model_status=mst, model_iters=mit, learned_status=lst, learned_iters=lit,
                n_per_round=res["n_per_round"][:p.iterations], field=field,
                counts=dict(blacklisted=int(black.sum()), failed=int(failed.sum()), training=len(train),
                            kept=len(res["kept"])))

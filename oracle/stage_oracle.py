"""CPU restatement of the whole per-star fitting stage, composed from the oracle's pieces in the order of
shopt/shopt.jl (:130-184 model fit + s-blacklist, :281-314 PCA pixel-grid fit, :357-417 learned fit + failed list,
:472-490 training split, :534-577 polyfit rounds).  TEST INFRASTRUCTURE ONLY (see shopt_oracle.h); PARITY UNPINNED.

Stars keep their original indices here (flags instead of splices), like the product's shopt_stage_run; the training
set is the first round(training_ratio * n_survivors) survivors of `perm`.
"""
import numpy as np

from . import shopt_oracle as orc

FLAG_BLACKLISTED, FLAG_PCA_FAILED, FLAG_LEARNED_OOB, FLAG_TRAINING, FLAG_REJECTED = 1, 2, 4, 8, 16


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


def run_stage(p, stamps, u, v, x0_model, x0_learned, perm, nthreads=1, fit_only=False, weights=None, x0_pixels=None):
    """p: any object with the shopt_params fields.  Returns the same dict layout as shopt_b200.pipeline.run_stage."""
    N = stamps.shape[0]
    s, g1, g2, _, _ = orc.fit_analytic_batch(stamps, p.n_a, x0_model, p.g_tol_model, p.max_iter, nthreads)
    model = np.stack([s, g1, g2])
    if fit_only:
        return dict(model=model)
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
    sl, g1l, g2l, _, _ = orc.fit_analytic_batch(pgf, p.n_a, x0_learned, p.g_tol_learned, p.max_iter, nthreads)
    flags[(sl < p.s_lo) | (sl > p.s_hi)] |= FLAG_LEARNED_OOB                        # shopt.jl:384-386
    surv = flags == 0
    field = field_interp(u[surv], v[surv], sl[surv], g1l[surv], g2l[surv])        # shopt.jl:422-466
    n_train = int(np.rint(p.training_ratio * surv.sum()))                            # round(Int, .): half to even
    train = np.array([g for g in perm if surv[g]][:n_train], dtype=np.int64)
    flags[train] |= FLAG_TRAINING
    res = orc.polyfit(pgf, u, v, train, p.degree, p.iterations, bool(p.median_constraint), bool(p.sigma_clip), nthreads)
    rejected = np.setdiff1d(train, res["kept"])
    flags[rejected] |= FLAG_REJECTED
    black = (flags & FLAG_BLACKLISTED) != 0
    failed = ~black & ((flags & (FLAG_PCA_FAILED | FLAG_LEARNED_OOB)) != 0)
    return dict(model=model, learned=np.stack([sl, g1l, g2l]), pgf=pgf, flags=flags, pca_k=k, poly=res["poly"],
                n_per_round=res["n_per_round"][:p.iterations], field=field,
                counts=dict(blacklisted=int(black.sum()), failed=int(failed.sum()), training=len(train),
                            kept=len(res["kept"])))

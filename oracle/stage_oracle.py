"""CPU restatement of the whole per-star fitting stage, composed from the oracle's pieces in the order of
shopt/shopt.jl (:130-184 model fit + s-blacklist, :281-314 PCA pixel-grid fit, :357-417 learned fit + failed list,
:472-490 training split, :534-577 polyfit rounds).  TEST INFRASTRUCTURE ONLY (see shopt_oracle.h); PARITY UNPINNED.

Stars keep their original indices here (flags instead of splices), like the product's shopt_stage_run; the training
set is the first round(training_ratio * n_survivors) survivors of `perm`.
"""
import numpy as np

from . import shopt_oracle as orc

FLAG_BLACKLISTED, FLAG_PCA_FAILED, FLAG_LEARNED_OOB, FLAG_TRAINING, FLAG_REJECTED = 1, 2, 4, 8, 16


def run_stage(p, stamps, u, v, x0_model, x0_learned, perm, nthreads=1, fit_only=False):
    """p: any object with the shopt_params fields.  Returns the same dict layout as shopt_b200.pipeline.run_stage."""
    N = stamps.shape[0]
    s, g1, g2, _, _ = orc.fit_analytic_batch(stamps, p.n_a, x0_model, p.g_tol_model, p.max_iter, nthreads)
    model = np.stack([s, g1, g2])
    if fit_only:
        return dict(model=model)
    flags = np.zeros(N, np.int32)
    flags[(s < p.s_lo) | (s > p.s_hi)] |= FLAG_BLACKLISTED                         # shopt.jl:166-171
    pgf, k, pst = orc.pixel_grid_pca_batch(stamps, p.pca_terms, p.pratio, p.lanczos, bool(p.unity_sum), nthreads)
    flags[pst == 1] |= FLAG_PCA_FAILED                                               # shopt.jl:305-309
    sl, g1l, g2l, _, _ = orc.fit_analytic_batch(pgf, p.n_a, x0_learned, p.g_tol_learned, p.max_iter, nthreads)
    flags[(sl < p.s_lo) | (sl > p.s_hi)] |= FLAG_LEARNED_OOB                        # shopt.jl:384-386
    surv = flags == 0
    n_train = int(np.rint(p.training_ratio * surv.sum()))                            # round(Int, .): half to even
    train = np.array([g for g in perm if surv[g]][:n_train], dtype=np.int64)
    flags[train] |= FLAG_TRAINING
    res = orc.polyfit(pgf, u, v, train, p.degree, p.iterations, bool(p.median_constraint), bool(p.sigma_clip), nthreads)
    rejected = np.setdiff1d(train, res["kept"])
    flags[rejected] |= FLAG_REJECTED
    black = (flags & FLAG_BLACKLISTED) != 0
    failed = ~black & ((flags & (FLAG_PCA_FAILED | FLAG_LEARNED_OOB)) != 0)
    return dict(model=model, learned=np.stack([sl, g1l, g2l]), pgf=pgf, flags=flags, pca_k=k, poly=res["poly"],
                n_per_round=res["n_per_round"][:p.iterations],
                counts=dict(blacklisted=int(black.sum()), failed=int(failed.sum()), training=len(train),
                            kept=len(res["kept"])))

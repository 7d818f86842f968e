"""Synthetic catalogues for the Kolmogorov-profile tests (independent numpy code: nothing here calls the oracle or the
library).  fkolmogorov of radialProfiles.jl:11-24 in closed form, f = (1 / 2 pi) C / (C^2 + kappa q)^(3/2)."""
import math

import numpy as np

C_KOLMO = (24 * math.gamma(6.5) * 0.2) ** (5 / 6) / (2 * math.pi) ** (5 / 3)


def np_e_to_g(e1, e2):
    ell = np.sqrt(e1 ** 2 + e2 ** 2)
    normG = np.sqrt(1 + 0.5 * (1 / ell ** 2 - np.sqrt(4 / ell ** 2 + 1 / ell ** 4)))
    return e1 * normG / ell, e2 * normG / ell


def np_profile(n, g1, g2, s, uc, vc):
    """(n, n) array indexed [v, u] (Julia column-major seen from numpy), 1-based pixel coordinates."""
    x = np.arange(1, n + 1, dtype=np.float64)
    du, dv = (x - uc)[None, :], (x - vc)[:, None]
    w1 = (1 + g1) * du + g2 * dv
    w2 = g2 * du + (1 - g1) * dv
    rho2 = (s * s / (1 - g1 * g1 - g2 * g2)) * (w1 * w1 + w2 * w2)
    return (1 / (2 * np.pi)) * C_KOLMO / (C_KOLMO ** 2 + rho2) ** 1.5


def np_cost(stamp, n_a, x):
    """analyticLBFGS.jl:8-46 with fkolmogorov in place of fGaussian; stamp (c, r) Julia layout."""
    c, r = stamp.shape
    g1, g2 = np_e_to_g(x[1], x[2])
    r0, c0 = (r - n_a) // 2, (c - n_a) // 2
    crop = stamp[c0:c0 + n_a, r0:r0 + n_a]
    f = np_profile(n_a, g1, g2, x[0] ** 2, n_a / 2.0, n_a / 2.0)
    res = f / f.sum() - crop
    return 0.5 * np.nansum(res * res)


def make_catalog(N, r, n_a, seed=11, noise=1e-6, frac_masked=0.0):
    """N unit-sum Kolmogorov stamps (N, r, r) centred like the analytic cost expects, truth s in [1.5, 3] (the profile's
    core radius C / s^2-ish pixels must sit inside the window for s to be measurable), |g| >= 0.03."""
    rng = np.random.default_rng(seed)
    u, v = rng.uniform(-1, 1, N), rng.uniform(-1, 1, N)
    s = 2.2 + 0.5 * u - 0.3 * v
    g1 = 0.08 + 0.03 * u - 0.02 * v * v
    g2 = -0.03 + 0.03 * v + 0.02 * u * u
    r0 = (r - n_a) // 2
    stamps = np.empty((N, r, r))
    for i in range(N):
        f = np_profile(r, g1[i], g2[i], s[i], n_a / 2.0 + r0, n_a / 2.0 + r0)
        stamps[i] = f / f.sum() + noise * rng.standard_normal((r, r))
    if frac_masked > 0:
        stamps[rng.uniform(size=stamps.shape) < frac_masked] = np.nan
    x0 = rng.uniform(0.6, 1.6, (N, 3)) * np.array([1.0, 0.5, 0.5])
    return dict(stamps=stamps, u=u, v=v, truth=np.stack([s, g1, g2], 1), x0_model=x0,
                x0_learned=rng.uniform(0.6, 1.6, (N, 3)) * np.array([1.0, 0.5, 0.5]), perm=rng.permutation(N).astype(np.int64))

"""Deterministic synthetic star catalogues for parity tests and bench.py.

Not part of the fitting path: it only manufactures inputs.  Works on any torch device so that the
100k-star bench catalogue is generated straight into HBM; tests call it on the CPU and hand the same
arrays to the oracle and to the CUDA path.

Layout: stamps is a (N, c, r) float64 tensor whose memory order is Julia's column-major (r, c, N):
pixel (u, v), 1-based, of star i is stamps[i, v-1, u-1].
"""
import math

import torch

SEED = 20240906


def truth_params(u, v):
    """Field-dependent truth: s in [0.20,0.40], |g| >= 0.03 everywhere."""
    s = 0.30 + 0.05 * u - 0.03 * v + 0.02 * u * v
    g1 = 0.08 + 0.03 * u - 0.02 * v * v
    g2 = -0.03 + 0.03 * v + 0.02 * u * u
    return s, g1, g2


def gaussian_stamps(s, g1, g2, r, c, uc, vc, device, dtype=torch.float64):
    """A * fGaussian(u, v, g1, g2, s, uc, vc) on the full r x c grid, unit sum
    (radialProfiles.jl:5-9 evaluated like analyticLBFGS.jl:24-34)."""
    uu = torch.arange(1, r + 1, device=device, dtype=dtype) - uc          # rows (fast axis)
    vv = torch.arange(1, c + 1, device=device, dtype=dtype) - vc
    du = uu.view(1, 1, r)
    dv = vv.view(1, c, 1)
    g1 = g1.view(-1, 1, 1); g2 = g2.view(-1, 1, 1); s = s.view(-1, 1, 1)
    w1 = (1 + g1) * du + g2 * dv
    w2 = g2 * du + (1 - g1) * dv
    ms = (s * s) / (1 - g1 * g1 - g2 * g2)
    f = torch.exp(-(ms * (w1 * w1 + w2 * w2)))
    return f / f.sum(dim=(1, 2), keepdim=True)


def make_catalog(N, r, n_a, seed=SEED, noise=1e-5, device="cpu", chunk=2048, frac_masked=0.0,
                 frac_contaminated=0.0, frac_bad_s=0.0, out=None):
    """Returns dict(stamps (N,r,r), u, v, x0_model (N,3), x0_learned (N,3), truth (N,3), perm (N,) int64).

    The profile centre is crop_start + n_a/2 (the centre the analytic model assumes inside the
    get_middle_nxn crop, analyticLBFGS.jl:20,28); for r == n_a this is r/2 as in SURVEY.md 8d.
    frac_masked: fraction of pixels set to NaN (what nanMask leaves behind, masks.jl:5-17);
    frac_contaminated: stars with a second Gaussian of 30 % flux at a random offset;
    frac_bad_s: stars whose truth s is outside [0.075, 1] (exercise shopt.jl:166-171).
    """
    dev = torch.device(device)
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    f64 = torch.float64
    u = torch.rand(N, generator=g, device=dev, dtype=f64) * 2 - 1
    v = torch.rand(N, generator=g, device=dev, dtype=f64) * 2 - 1
    x0_model = torch.rand(N, 3, generator=g, device=dev, dtype=f64)
    x0_learned = torch.rand(N, 3, generator=g, device=dev, dtype=f64)
    perm = torch.argsort(torch.rand(N, generator=g, device=dev, dtype=f64))
    s, g1, g2 = truth_params(u, v)
    if frac_bad_s > 0:
        bad = torch.rand(N, generator=g, device=dev, dtype=f64) < frac_bad_s
        s = torch.where(bad, torch.full_like(s, 1.25), s)
    centre = (r - n_a) // 2 + n_a / 2.0
    stamps = out if out is not None else torch.empty(N, r, r, device=dev, dtype=f64)
    for a in range(0, N, chunk):
        b = min(N, a + chunk)
        st = gaussian_stamps(s[a:b], g1[a:b], g2[a:b], r, r, centre, centre, dev)
        if frac_contaminated > 0:
            cont = torch.rand(b - a, generator=g, device=dev, dtype=f64) < frac_contaminated
            off = (torch.rand(b - a, 2, generator=g, device=dev, dtype=f64) - 0.5) * 12.0
            st2 = torch.zeros_like(st)
            idx = torch.nonzero(cont).flatten()
            for j in idx.tolist():
                st2[j] = gaussian_stamps(s[a + j:a + j + 1], g1[a + j:a + j + 1], g2[a + j:a + j + 1], r, r,
                                         centre + float(off[j, 0]), centre + float(off[j, 1]), dev)[0]
            st = torch.where(cont.view(-1, 1, 1), 0.7 * st + 0.3 * st2, st)
        if noise > 0:
            st = st + noise * torch.randn(b - a, r, r, generator=g, device=dev, dtype=f64)
        if frac_masked > 0:
            m = torch.rand(b - a, r, r, generator=g, device=dev, dtype=f64) < frac_masked
            st = torch.where(m, torch.full_like(st, math.nan), st)
        stamps[a:b] = st
    return dict(stamps=stamps, u=u, v=v, x0_model=x0_model, x0_learned=x0_learned,
                truth=torch.stack([s, g1, g2], dim=1), perm=perm)

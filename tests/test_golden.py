"""Golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py from independent numpy/scipy code) against
the oracle on CPU and against the CUDA path through the C ABI on the GPU."""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name))


def check_analytic(cost_fn, grad_fn, fit_fn):
    d = load("analytic.npz")
    st, x0, n_a = d["stamps"], d["x0"], int(d["n_a"])
    f = cost_fn(st, n_a, x0)
    assert np.allclose(f, d["cost"], rtol=1e-10)
    g = grad_fn(st, n_a, x0)
    assert np.allclose(g, d["grad_fd"], rtol=5e-5, atol=1e-9)
    s, g1, g2 = fit_fn(st, n_a, x0)
    # the fit stops at |g|_inf <= 1e-8: within ~1e-6 of the true minimiser found by scipy
    assert np.allclose(s, d["minimum"][:, 0], rtol=2e-5)
    assert np.allclose(g1, d["minimum"][:, 1], atol=2e-5) and np.allclose(g2, d["minimum"][:, 2], atol=2e-5)


def check_pca(pca_fn):
    d = load("pca.npz")
    rec, k = pca_fn(d["stamps"])
    assert np.array_equal(k, d["k"])
    scale = np.abs(d["recon"]).max(axis=(1, 2), keepdims=True)
    assert np.max(np.abs(rec - d["recon"]) / scale) < 1e-9


def check_polyfit(polyfit_fn):
    d = load("polyfit.npz")
    res = polyfit_fn(d["pgf"], d["u"], d["v"], d["train"], int(d["degree"]), 3)
    assert res["n_per_round"].tolist() == d["sizes"].tolist()
    assert res["kept"].tolist() == d["kept"].tolist()
    assert np.allclose(res["poly"], d["poly"], rtol=1e-7, atol=1e-9)
    assert np.allclose(res["mse"], d["mse"], rtol=1e-8)


def test_oracle_matches_golden(orc):
    check_analytic(lambda st, n_a, x: np.array([orc.cost(st[i], n_a, x[i]) for i in range(len(st))]),
                   lambda st, n_a, x: np.array([orc.grad(st[i], n_a, x[i]) for i in range(len(st))]),
                   lambda st, n_a, x: orc.fit_analytic_batch(st, n_a, x, 1e-8)[:3])

    def pca(stamps):
        r = [orc.pca_image(orc.nan_to_zero(s), 50) for s in stamps]
        return np.stack([x[0] for x in r]), np.array([x[1] for x in r])
    check_pca(pca)
    check_polyfit(lambda *a: orc.polyfit(*a))


@pytest.mark.gpu
def test_cuda_path_matches_golden(stage):
    check_analytic(lambda st, n_a, x: stage.analytic_cost_grad_batch(st, n_a, x)[0],
                   lambda st, n_a, x: stage.analytic_cost_grad_batch(st, n_a, x)[1],
                   lambda st, n_a, x: stage.fit_analytic_batch(st, n_a, x, 1e-8)[:3])

    def pca(stamps):
        pgf, k, status = stage.pca_batch(stamps, 50, 0.99, 0, False)
        assert np.all(status == 0)
        return pgf, k
    check_pca(pca)
    check_polyfit(lambda *a: stage.polyfit(*a))

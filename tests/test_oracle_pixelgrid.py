"""CPU: the oracle's nanToZero / pca_image / smooth (masks.jl:38-50, pca.jl:1-18, dataPreprocessing.jl:94-139)
against LAPACK (numpy.linalg.svd) and analytic known-answer cases."""
import numpy as np
import pytest

from shopt_b200 import synth


def np_pca_image(img, ncomp, pratio=0.99):
    """MultivariateStats 0.10.2 fit(PCA; maxoutdim, pratio) / transform / reconstruct with LAPACK.
    img (c, r) Julia layout -> X = img.T is r x c, rows = variables, columns = observations."""
    X = img.T
    d, n = X.shape
    mu = X.mean(axis=1, keepdims=True)
    Z = X - mu
    U, S, _ = np.linalg.svd(Z, full_matrices=False)
    var = S ** 2 / (n - 1)
    k, acc = 1, var[0]
    while k < len(var) and k < ncomp and acc < pratio * var.sum():
        acc += var[k]
        k += 1
    P = U[:, :k]
    return (P @ (P.T @ Z) + mu).T, k


def test_nan_to_zero(orc):
    a = np.array([[1.0, np.nan], [np.nan, -2.0]])
    assert np.array_equal(orc.nan_to_zero(a), np.array([[1.0, 0.001], [0.001, -2.0]]))   # masks.jl:43


@pytest.mark.parametrize("r,c,ncomp", [(24, 24, 50), (17, 17, 3), (20, 14, 50), (14, 20, 5)])
def test_pca_matches_lapack(orc, r, c, ncomp):
    rng = np.random.default_rng(r * 100 + c)
    # a few strong components + a noise floor so that the 99 % rule picks a well-separated k
    img = sum(np.outer(rng.normal(size=c), rng.normal(size=r)) * w for w in (5.0, 2.0, 1.0)) + 1e-3 * rng.normal(size=(c, r))
    out, k, st = orc.pca_image(img, ncomp)
    want, kw = np_pca_image(img, ncomp)
    assert st == 0 and k == kw
    assert np.allclose(out, want, rtol=1e-9, atol=1e-11)


def test_pca_rank_k_is_reproduced_exactly(orc):
    rng = np.random.default_rng(3)
    img = np.outer(rng.normal(size=19), rng.normal(size=19)) + 0.5 * np.outer(rng.normal(size=19), rng.normal(size=19))
    out, k, st = orc.pca_image(img, 50, pratio=1.0 - 1e-13)
    assert st == 0 and k <= 3
    assert np.allclose(out, img, atol=1e-11)


def test_pca_on_star_stamps(orc):
    cat = synth.make_catalog(3, 31, 31)
    for img in cat["stamps"].numpy():
        out, k, st = orc.pca_image(img, 50)
        want, kw = np_pca_image(img, 50)
        assert st == 0 and k == kw
        assert np.allclose(out, want, rtol=1e-8, atol=1e-12)


def test_pca_nonfinite_input_fails(orc):
    img = np.ones((8, 8))
    img[2, 3] = np.inf
    out, k, st = orc.pca_image(img, 5)
    assert st == 1


def test_lanczos_kernel_is_a_delta(orc):
    # dataPreprocessing.jl:94-110: at integer offsets sinc(pi k) ~ 1e-16, so the normalised kernel is delta + O(1e-16)
    k = orc.lanczos_kernel2d(3)
    assert k.shape == (7, 7) and np.isclose(k.sum(), 1.0, rtol=1e-15)
    d = np.zeros((7, 7))
    d[3, 3] = 1.0
    assert np.max(np.abs(k - d)) < 1e-15


def test_smooth_identity_and_nan(orc):
    rng = np.random.default_rng(5)
    img = rng.normal(size=(12, 15))
    img[4, 7] = np.nan
    out = orc.smooth(img, 3)
    assert np.isnan(out[4, 7])                      # NaN pixels stay NaN (dataPreprocessing.jl:118)
    m = ~np.isnan(img)
    assert np.allclose(out[m], img[m], rtol=1e-13, atol=1e-14)


def test_pixel_grid_pca_batch_modes(orc):
    cat = synth.make_catalog(4, 21, 21, frac_masked=0.02)
    st = cat["stamps"].numpy()
    st[3, 0, 0] = np.inf                            # LAPACK would raise -> zeros + failed (shopt.jl:305-309)
    pgf, k, status = orc.pixel_grid_pca_batch(st, 50, 0.99, 3, unity_sum=False)
    assert status.tolist() == [0, 0, 0, 1] and np.all(pgf[3] == 0)
    for i in range(3):
        want, kw = np_pca_image(orc.nan_to_zero(st[i]), 50)
        assert k[i] == kw and np.allclose(pgf[i], want, rtol=1e-8, atol=1e-12)
    pgf1, _, _ = orc.pixel_grid_pca_batch(st, 50, 0.99, 3, unity_sum=True)
    for i in range(3):
        assert np.isclose(pgf1[i].sum(), 1.0, rtol=1e-12)          # shopt.jl:295


def test_chisq_mode_closed_form_is_what_lbfgs_converges_to():
    """chisq.jl:1-11 minimised by a quasi-Newton method from a random start (scipy's L-BFGS here, Optim's in the
    reference): pixels with a finite convex term end on the star, every other pixel keeps its start value."""
    from scipy.optimize import minimize
    from oracle import stage_oracle
    rng = np.random.default_rng(2)
    r = 6
    star = rng.normal(size=(r, r)) + 1.0
    w = rng.uniform(0.5, 2.0, size=(r, r))
    star[1, 2] = np.nan
    w[0, 0], w[3, 3], w[4, 1] = 0.0, np.inf, np.nan
    x0 = rng.uniform(size=r * r)

    def cost(p):
        with np.errstate(divide="ignore", invalid="ignore"):
            t = (p - star.ravel()) ** 2 / w.ravel()
        return np.where(np.isfinite(t), t, 0.0).sum()

    def grad(p):
        with np.errstate(divide="ignore", invalid="ignore"):
            t = (p - star.ravel()) ** 2 / w.ravel()
            g = 2.0 * (p - star.ravel()) / w.ravel()
        return np.where(np.isfinite(t) & np.isfinite(g), g, 0.0)

    res = minimize(cost, x0, jac=grad, method="L-BFGS-B", options=dict(gtol=1e-12, ftol=1e-30, maxiter=500))
    want = stage_oracle.pixel_grid_elementwise(2, star[None], 3, False, w[None], x0.reshape(1, r, r))[0]
    assert np.allclose(res.x.reshape(r, r), want, atol=1e-9)
    dead = ~(np.isfinite(star) & np.isfinite(w) & (w > 0))
    assert dead.sum() == 4 and np.array_equal(want[dead], x0.reshape(r, r)[dead])

"""GPU parity: PCA-mode pixel-grid fit (pixelgrid.cu) through the C ABI against the CPU oracle.
Tolerance: reconstructed pixels within 1e-6 relative to the stamp's peak (they feed the learned fit and the polynomial
coefficients, which carry the 1e-6 bar); number of components kept and failed-star flags bit-exact."""
import time

import numpy as np
import pytest

from shopt_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["fast", "one-stage", "direct", "jacobi"], autouse=True)
def pca_path(request, monkeypatch):
    """Every case runs four times: as shipped (pca_power_kernel first, pca_kernel on what it leaves: power path, block
    subspace iteration, direct Gram-matrix eigen path, Jacobi SVD when a check fails), with stage 1 off (pca_kernel's own
    power / block paths serve the one-component stars), with the fast paths off (direct path first for every star), and
    with the full Jacobi SVD forced for every star."""
    for name in ("SHOPT_PCA_NO_FAST", "SHOPT_PCA_NO_EIG", "SHOPT_PCA_DEBUG", "SHOPT_PCA_NO_LIGHT"):
        monkeypatch.delenv(name, raising=False)
    if request.param == "one-stage":
        monkeypatch.setenv("SHOPT_PCA_NO_LIGHT", "1")
    if request.param in ("direct", "jacobi"):
        monkeypatch.setenv("SHOPT_PCA_NO_FAST", "1")
    if request.param == "jacobi":
        monkeypatch.setenv("SHOPT_PCA_NO_EIG", "1")
    return request.param


def close(a, b, tol=1e-6):
    scale = np.abs(b).max(axis=(1, 2), keepdims=True)
    scale[scale == 0] = 1.0
    return np.max(np.abs(a - b) / scale) <= tol


@pytest.mark.parametrize("N,r,kw", [(24, 131, {}), (16, 75, {}), (12, 64, dict(frac_masked=0.01)), (12, 33, {}),
                                    (8, 131, dict(noise=1e-3)), (8, 100, dict(noise=3e-4, frac_contaminated=0.5))])
def test_pca_star_stamps(stage, orc, N, r, kw):
    cat = synth.make_catalog(N, r, min(r, 64), **kw)
    st = cat["stamps"].numpy()
    for unity in (False, True):
        pgf, k, status = stage.pca_batch(st, 50, 0.99, 3, unity)
        want, kw_, st_ = orc.pixel_grid_pca_batch(st, 50, 0.99, 3, unity, nthreads=orc.max_threads())
        assert np.array_equal(status, st_) and np.array_equal(k, kw_)
        assert close(pgf, want)


def test_few_masked_pixels_stay_one_component(stage, orc):
    """A handful of masked pixels (NaN -> 0.001) does not add components: stage 1 serves these stars through its
    NaN-aware pass; and a stamp with an Inf is a failed star whichever kernel meets it first."""
    cat = synth.make_catalog(20, 131, 64)
    st = cat["stamps"].numpy().copy()
    rng = np.random.default_rng(3)
    for i in range(20):
        for _ in range(1 + i % 3):
            st[i, rng.integers(0, 131), rng.integers(0, 131)] = np.nan
    st[7, 5, 9] = np.inf
    for unity in (False, True):
        pgf, k, status = stage.pca_batch(st, 50, 0.99, 3, unity)
        want, kw_, st_ = orc.pixel_grid_pca_batch(st, 50, 0.99, 3, unity, nthreads=orc.max_threads())
        assert np.array_equal(status, st_) and np.array_equal(k, kw_) and status[7] == 1
        assert (k[np.arange(20) != 7] == 1).all()
        assert close(pgf, want)


@pytest.mark.parametrize("r,c,ncomp", [(24, 24, 50), (17, 17, 3), (40, 28, 50), (28, 40, 5), (131, 131, 50), (2, 2, 1)])
def test_pca_generic_matrices(stage, orc, r, c, ncomp):
    rng = np.random.default_rng(r * 100 + c)
    N = 6
    img = np.zeros((N, c, r))
    for i in range(N):
        nc = 1 + i
        img[i] = sum(np.outer(rng.normal(size=c), rng.normal(size=r)) * (3.0 / (1 + j)) for j in range(nc))
        img[i] += 1e-3 * rng.normal(size=(c, r))
    pgf, k, status = stage.pca_batch(img, ncomp, 0.99, 0, False)       # lanczos_a = 0: pca_image alone
    res = [orc.pca_image(img[i], ncomp, 0.99) for i in range(N)]
    assert np.all(status == 0) and k.tolist() == [x[1] for x in res]
    assert close(pgf, np.stack([x[0] for x in res]), 1e-9)


def test_noise_dominated_stamps_keep_many_components(stage, orc):
    rng = np.random.default_rng(9)
    img = rng.normal(size=(4, 60, 60))
    pgf, k, status = stage.pca_batch(img, 50, 0.99, 3, False)
    want, kw_, st_ = orc.pixel_grid_pca_batch(img, 50, 0.99, 3, False, nthreads=4)
    assert np.all(k >= 40) and np.array_equal(k, kw_) and np.all(status == 0)
    assert close(pgf, want, 1e-8)


def test_flat_spectrum_stamps_take_the_direct_path(stage, orc, pca_path, monkeypatch):
    """Masked / noisy 131 x 131 stamps keep tens of components: the direct path (tridiagonalised Gram matrix, Sturm
    multisection, inverse iteration) must serve them, and agree with the oracle's SVD."""
    rng = np.random.default_rng(21)
    r = 131
    yy, xx = np.mgrid[:r, :r]
    img = np.empty((6, r, r))
    for i in range(6):
        g = np.exp(-((yy - 65.0) ** 2 + (xx - 65.0) ** 2) / (2 * (2.0 + 0.3 * i) ** 2))
        img[i] = g / g.sum() + (1e-4 * (i + 1)) * rng.normal(size=(r, r))
    img[2][rng.random((r, r)) < 0.01] = np.nan
    monkeypatch.setenv("SHOPT_PCA_DEBUG", "1")               # status >> 8: 100 direct path, 200 Jacobi after a failed check
    pgf, k, status = stage.pca_batch(img, 50, 0.99, 3, False)
    want, kw_, st_ = orc.pixel_grid_pca_batch(img, 50, 0.99, 3, False, nthreads=orc.max_threads())
    assert np.array_equal(k, kw_) and np.all((status & 0xff) == 0)
    assert close(pgf, want, 1e-9)
    assert k.max() > 6
    if pca_path != "jacobi":
        assert np.all((status >> 8)[k > 6] == 100), (status >> 8, k)


def test_failed_stars_and_exact_rank(stage, orc):
    cat = synth.make_catalog(5, 41, 32)
    st = cat["stamps"].numpy()
    st[1, 3, 4] = np.inf                      # LAPACK raises -> zeros(r,c) + failedStars (shopt.jl:305-309)
    st[3, 7, 7] = -np.inf
    st[4] = np.nan                            # all NaN -> nanToZero makes it constant 0.001
    pgf, k, status = stage.pca_batch(st, 50, 0.99, 3, False)
    want, kw_, st_ = orc.pixel_grid_pca_batch(st, 50, 0.99, 3, False)
    assert status.tolist() == st_.tolist() == [0, 1, 0, 1, 0]
    assert np.all(pgf[1] == 0) and np.all(pgf[3] == 0)
    assert np.array_equal(k, kw_)
    assert close(pgf, want)
    rng = np.random.default_rng(3)
    img = (np.outer(rng.normal(size=19), rng.normal(size=19)) + 0.5 * np.outer(rng.normal(size=19), rng.normal(size=19)))[None]
    pgf, k, status = stage.pca_batch(img, 50, 1.0 - 1e-13, 0, False)
    assert k[0] <= 3 and np.allclose(pgf, img, atol=1e-11)            # rank-k stamps are reproduced by k components


def test_smooth_batch(stage, orc):
    rng = np.random.default_rng(5)
    img = rng.normal(size=(3, 15, 12))
    img[0, 4, 7] = np.nan
    img[1, 0, 0] = np.nan
    for nz in (False, True):
        got = stage.smooth_batch(img, 3, nz)
        for i in range(3):
            src = orc.nan_to_zero(img[i]) if nz else img[i]
            want = orc.smooth(src, 3)
            assert np.array_equal(np.isnan(got[i]), np.isnan(want))
            assert np.allclose(got[i], want, rtol=1e-13, atol=1e-15, equal_nan=True)


def test_oversampled_stamps_global_scratch_path(stage, orc):
    """301 x 301 (BASELINE config 5) does not fit shared memory: same kernel, W in an L2-resident scratch block."""
    cat = synth.make_catalog(2, 301, 128)
    st = cat["stamps"].numpy()
    pgf, k, status = stage.pca_batch(st, 50, 0.99, 3, False)
    want, kw_, st_ = orc.pixel_grid_pca_batch(st, 50, 0.99, 3, False, nthreads=2)
    assert np.array_equal(k, kw_) and np.all(status == 0)
    assert close(pgf, want)

"""GPU parity: ingest kernels (nanMask, unity sum, crop / bilinear resample + renormalise + smooth) against the numpy
restatement of cataloging() (dataPreprocessing.jl:342-397).  Tolerance 1e-12 of the stamp peak; NaN masks bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def raw_vignets(seed, N, r, dtype):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:r, 0:r]
    img = np.stack([np.exp(-((x - r / 2 + rng.normal()) ** 2 + (y - r / 2 + rng.normal()) ** 2) / (2 * (2 + rng.uniform()) ** 2))
                    for _ in range(N)]) * 1000 + rng.normal(size=(N, r, r))
    img[rng.uniform(size=img.shape) < 0.01] = -1e30          # SExtractor's masked-pixel sentinel
    return img.astype(dtype)


def close(a, b, tol=1e-12):
    scale = np.nanmax(np.abs(b), axis=(1, 2), keepdims=True)
    return np.array_equal(np.isnan(a), np.isnan(b)) and np.nanmax(np.abs(a - b) / scale) <= tol


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("unity", [False, True])
@pytest.mark.parametrize("r0,new_dim", [(25, None), (25, 15), (25, 16), (24, 17), (15, 33), (16, 40)])
def test_ingest_matches_cataloging(stage, orc, dtype, unity, r0, new_dim):
    from oracle import stage_oracle
    v = raw_vignets(r0 * 3 + (new_dim or 0), 5, r0, dtype)
    got = stage.ingest_batch(v, unity, new_dim, 3)
    want = stage_oracle.ingest(v, unity, new_dim, 3)
    assert got.shape == want.shape and close(got, want)
    if new_dim is not None:
        assert np.allclose(np.nansum(orc.nan_to_zero(got[0])), 1.0, atol=0.05)     # renormalised (B.9), then smoothed


def test_ingest_feeds_the_stage(stage, orc):
    """float32 vignets -> ingest -> analytic fit: the chain the Julia shim calls."""
    v = raw_vignets(1, 8, 41, np.float32)
    stamps = stage.ingest_batch(v, True)
    x0 = np.full((8, 3), 0.4)
    s, g1, g2, status, iters = stage.fit_analytic_batch(stamps, 32, x0)[:5]
    assert np.all(np.isfinite(s)) and np.all(s > 0)

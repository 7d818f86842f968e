"""GPU parity: per-pixel polynomial interpolation (polyfit.cu) through the C ABI against the CPU oracle.
Tolerance (north_star): polynomial coefficients within 1e-6 relative; outlier selection and star indexing bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make_case(seed, N, r, c, degree, noise=0.01, n_out=6, n_train=None):
    rng = np.random.default_rng(seed)
    u, v = rng.uniform(-1, 1, N), rng.uniform(-1, 1, N)
    K = (degree + 1) * (degree + 2) // 2
    terms = [(p, q) for p in range(degree + 1) for q in range(degree + 1 - p)]
    V = np.stack([u ** p * v ** q for p, q in terms], axis=1)
    truth = rng.normal(size=(K, c, r)) + 2.0
    pgf = np.einsum("nk,kcr->ncr", V, truth) + noise * rng.normal(size=(N, c, r))
    if n_out:
        pgf[rng.choice(N, n_out, replace=False)] += rng.uniform(0.5, 2.0, (n_out, 1, 1))
    train = rng.permutation(N)[:n_train if n_train is not None else int(0.8 * N)]
    return pgf, u, v, train, truth


def coef_close(a, b, rtol=1e-6):
    scale = np.abs(b).max(axis=(1, 2), keepdims=True)       # per-term scale: coefficients of a term share a magnitude
    return np.max(np.abs(a - b) / scale) <= rtol


@pytest.mark.parametrize("degree,N,r,c,iterations", [(1, 300, 9, 7, 3), (3, 500, 13, 13, 3), (2, 200, 5, 40, 1),
                                                     (5, 600, 11, 11, 3), (4, 300, 8, 8, 2), (6, 500, 6, 6, 2),
                                                     (3, 2000, 33, 31, 3)])
def test_polyfit_matches_oracle(stage, orc, degree, N, r, c, iterations):
    pgf, u, v, train, _ = make_case(degree * 7 + N, N, r, c, degree)
    got = stage.polyfit(pgf, u, v, train, degree, iterations)
    want = orc.polyfit(pgf, u, v, train, degree, iterations, nthreads=orc.max_threads())
    assert got["n_per_round"].tolist() == want["n_per_round"].tolist()
    assert got["kept"].tolist() == want["kept"].tolist()            # outlier selection bit-exact
    assert coef_close(got["poly"], want["poly"])
    assert np.allclose(got["mse"], want["mse"], rtol=1e-9)


def test_sigma_clip_variant_and_exact_polynomial(stage, orc):
    pgf, u, v, train, truth = make_case(5, 400, 10, 10, 3, noise=0.0, n_out=0)
    got = stage.polyfit(pgf, u, v, train, 3, 2, sigma_clip=True)
    assert coef_close(got["poly"], truth, 1e-9)                      # pixels that ARE polynomials give them back
    pgf, u, v, train, _ = make_case(6, 400, 10, 10, 2, n_out=5)
    got = stage.polyfit(pgf, u, v, train, 2, 3, sigma_clip=True)
    want = orc.polyfit(pgf, u, v, train, 2, 3, sigma_clip=True)
    assert got["kept"].tolist() == want["kept"].tolist() and coef_close(got["poly"], want["poly"])


def test_small_training_sets_and_zero_rounds(stage, orc):
    pgf, u, v, _, _ = make_case(8, 30, 6, 6, 1)
    for n in (0, 1, 2, 3, 4):                                        # interpolate.jl:142-145: [median, 0, 0]
        train = np.arange(n) * 2
        got = stage.polyfit(pgf, u, v, train, 1, 1)
        if n == 0:
            continue
        want = orc.polyfit(pgf, u, v, train, 1, 1)
        assert np.allclose(got["poly"], want["poly"], rtol=1e-15, atol=0)
        assert np.allclose(got["mse"], want["mse"], rtol=1e-12)
    got = stage.polyfit(pgf, u, v, np.arange(20), 1, 0)
    assert np.all(got["poly"] == 1.0)                                # ones(r, c, K), shopt.jl:492
    got = stage.polyfit(pgf, u, v, np.arange(12), 1, 3)              # 12 -> 10 -> 9 stars
    want = orc.polyfit(pgf, u, v, np.arange(12), 1, 3)
    assert got["n_per_round"].tolist() == want["n_per_round"].tolist() == [12, 10, 9]
    assert got["kept"].tolist() == want["kept"].tolist()


def test_psf_at_batch(stage, orc):
    rng = np.random.default_rng(4)
    for degree, r, c in [(1, 9, 9), (3, 31, 29), (5, 12, 12)]:
        K = (degree + 1) * (degree + 2) // 2
        poly = rng.normal(size=(K, c, r)) * 0.1
        poly[0] += 1.0
        u, v = rng.uniform(-1, 1, 17), rng.uniform(-1, 1, 17)
        got = stage.psf_at_batch(poly, degree, u, v)
        for m in range(17):
            want = orc.psf_at(poly, degree, u[m], v[m])
            assert np.allclose(got[m], want, rtol=1e-10, atol=1e-15)
            assert np.isclose(got[m].sum(), 1.0, rtol=1e-12)


def test_building_blocks_shard_additively(stage):
    """Multi-GPU contract: the [V'V | V'P] accumulators of disjoint star shards add up to the accumulator of the
    union (what the NCCL all-reduce relies on), and linearity in the pixel-grid fits holds."""
    import ctypes as C
    import torch
    from shopt_b200 import _lib
    L = _lib.lib()
    N, r, c, deg = 900, 21, 19, 3
    g = torch.Generator(device="cuda").manual_seed(1)
    pgf = torch.rand(N, c, r, generator=g, device="cuda", dtype=torch.float64)
    u = torch.rand(N, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    v = torch.rand(N, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    n_acc = L.shopt_polyfit_acc_len(r, c, deg)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def acc_of(idx):
        idx = idx.contiguous()
        acc = torch.empty(n_acc, device="cuda", dtype=torch.float64)
        _lib.check(L.shopt_polyfit_normal_dev(pgf.data_ptr(), r, c, u.data_ptr(), v.data_ptr(), idx.data_ptr(), len(idx),
                                              deg, None, acc.data_ptr(), st))
        torch.cuda.synchronize()
        return acc
    idx = torch.randperm(N, device="cuda")[:700]
    whole = acc_of(idx)
    parts = acc_of(idx[:300]) + acc_of(idx[300:450]) + acc_of(idx[450:])
    assert torch.allclose(whole, parts, rtol=1e-12, atol=1e-12)
    assert torch.equal(acc_of(idx), whole)                            # deterministic: no floating-point atomics


@pytest.mark.parametrize("degree,N,r,c", [(3, 1500, 37, 35), (5, 900, 41, 43), (1, 700, 9, 7), (7, 600, 17, 19)])
def test_ring_kernels_equal_the_plain_load_kernels_bit_for_bit(monkeypatch, degree, N, r, c):
    """normal_ring_kernel / mse_ring_kernel (per-lane cp.async rings, the default) perform the arithmetic of normal_kernel /
    mse_kernel in the same order: [V'V | V'P] and the per-star errors are bit-identical with SHOPT_POLY_RING=0 and 3, on
    list lengths that are no multiple of the 64-star refill or of the ring depth, ragged pixel tiles and NaN values.
    Pixels per thread (SHOPT_POLY_PX4, 10 < K <= 21) regroup the star chunks and the per-star pixel sums: rounding only."""
    import ctypes as C
    import torch
    from shopt_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(degree)
    pgf = torch.rand(N, c, r, generator=g, device="cuda", dtype=torch.float64)
    pgf[5, 3, 2] = float("nan")
    u = torch.rand(N, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    v = torch.rand(N, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    K = (degree + 1) * (degree + 2) // 2
    poly = torch.rand(K, c, r, generator=g, device="cuda", dtype=torch.float64)
    n_acc = L.shopt_polyfit_acc_len(r, c, degree)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run(idx, ring, px4="1"):
        monkeypatch.setenv("SHOPT_POLY_RING", ring)
        monkeypatch.setenv("SHOPT_POLY_PX4", px4)
        acc = torch.empty(n_acc, device="cuda", dtype=torch.float64)
        mse = torch.empty(len(idx), device="cuda", dtype=torch.float64)
        _lib.check(L.shopt_polyfit_normal_dev(pgf.data_ptr(), r, c, u.data_ptr(), v.data_ptr(), idx.data_ptr(), len(idx),
                                              degree, None, acc.data_ptr(), st))
        _lib.check(L.shopt_polyfit_mse_dev(pgf.data_ptr(), r, c, u.data_ptr(), v.data_ptr(), idx.data_ptr(), len(idx),
                                           degree, poly.data_ptr(), mse.data_ptr(), st))
        torch.cuda.synchronize()
        return acc.cpu().numpy(), mse.cpu().numpy()

    for n in (N - 3, 131, 64, 9, 1):
        idx = torch.randperm(N, generator=g, device="cuda")[:n].contiguous()
        acc_r, mse_r = run(idx, "3", px4="0")                                    # same tiling in both forms
        acc_p, mse_p = run(idx, "0", px4="0")
        assert np.array_equal(acc_r, acc_p, equal_nan=True)
        assert np.array_equal(mse_r, mse_p, equal_nan=True)
        acc_4, mse_4 = run(idx, "3", px4="1")                                    # the default
        assert np.allclose(acc_4, acc_r, rtol=1e-12, atol=1e-12, equal_nan=True)     # other tile: other star chunks
        assert np.allclose(mse_4, mse_r, rtol=1e-13, atol=0, equal_nan=True)


@pytest.mark.parametrize("degree,N,r,c", [(1, 301, 9, 7), (3, 500, 13, 13), (2, 1000, 21, 19)])
def test_median_constraint(stage, orc, degree, N, r, c):
    """interpolate.jl:118-124, 139-141: coefficient 1 = per-pixel median over the training stars (radix select on the
    device); odd and even training-set sizes, ties."""
    pgf, u, v, train, _ = make_case(degree * 5 + N, N, r, c, degree)
    pgf[:, 0, 0] = np.round(pgf[:, 0, 0], 1)                        # many exact ties in one pixel
    pgf[:, 1, 1] = -np.abs(pgf[:, 1, 1])                            # negative values exercise the key ordering
    pgf[::2, 2, 2] *= -1.0                                           # mixed signs
    got = stage.polyfit(pgf, u, v, train, degree, 3, median_constraint=True)
    want = orc.polyfit(pgf, u, v, train, degree, 3, median_constraint=True, nthreads=orc.max_threads())
    assert got["n_per_round"].tolist() == want["n_per_round"].tolist()
    assert got["kept"].tolist() == want["kept"].tolist()
    assert np.array_equal(got["poly"][0], want["poly"][0])          # the median itself is exact
    assert coef_close(got["poly"], want["poly"])
    assert np.allclose(got["mse"], want["mse"], rtol=1e-9)


def test_nan_pixel_grid_values_are_replaced_by_the_median(stage, orc):
    """interpolate.jl:126-130 (median_constraint = false): NaN values of a pixel become the median of its finite
    values before the solve; the stars keep their NaNs, so their reconstruction error is NaN and they sort first."""
    pgf, u, v, train, _ = make_case(3, 400, 9, 8, 2, n_out=0)
    rng = np.random.default_rng(1)
    for _ in range(25):
        pgf[rng.integers(400), rng.integers(8), rng.integers(9)] = np.nan
    pgf[train[:7], 3, 4] = np.nan
    got = stage.polyfit(pgf, u, v, train, 2, 2)
    want = orc.polyfit(pgf.copy(), u, v, train, 2, 2, nthreads=orc.max_threads())
    assert got["n_per_round"].tolist() == want["n_per_round"].tolist()
    assert got["kept"].tolist() == want["kept"].tolist()
    assert np.all(np.isfinite(got["poly"])) and coef_close(got["poly"], want["poly"])
    assert np.array_equal(np.isnan(got["mse"]), np.isnan(want["mse"]))
    assert np.allclose(got["mse"], want["mse"], rtol=1e-9, equal_nan=True)


def _design(u, v, degree):
    return np.stack([u ** p * v ** q for p in range(degree + 1) for q in range(degree + 1 - p)], axis=1)


@pytest.mark.parametrize("degree", [1, 2, 3])
def test_raw_sky_coordinates(stage, orc, degree):
    """The reference hands RAW sky coordinates to polynomial_optimizer (shopt.jl:472-490 pass u_coordinates = RA,
    v_coordinates = Dec in degrees): u ~ 150, v ~ 2, field 0.2 deg wide, cond(A) up to 1e17.  Julia's A \\ z (pivoted QR,
    rank truncation, minimum norm) still fits; the product's centred-basis solve must give finite coefficients whose
    FITTED VALUES agree with the oracle's QR to 1e-6, with the same rank decision and the same rejected stars."""
    rng = np.random.default_rng(40 + degree)
    N, r, c = 1500, 9, 8
    a, b = rng.uniform(-1, 1, N), rng.uniform(-1, 1, N)
    u, v = 150.0 + 0.1 * a, 2.2 + 0.1 * b
    K = (degree + 1) * (degree + 2) // 2
    truth = rng.normal(size=(K, c, r)) * 0.1 + 1.0
    pgf = np.einsum("nk,kcr->ncr", _design(a, b, degree), truth) + 1e-3 * rng.normal(size=(N, c, r))
    pgf[rng.choice(N, 8, replace=False)] += 0.5
    train = rng.permutation(N)[:1200]
    got = stage.polyfit(pgf, u, v, train, degree, 3)
    want = orc.polyfit(pgf, u, v, train, degree, 3, nthreads=orc.max_threads())
    assert np.all(np.isfinite(got["poly"]))
    assert got["n_per_round"].tolist() == want["n_per_round"].tolist()
    assert got["kept"].tolist() == want["kept"].tolist()
    V = _design(u[got["kept"]], v[got["kept"]], degree)
    fit_g = np.einsum("nk,kcr->ncr", V, got["poly"])
    fit_w = np.einsum("nk,kcr->ncr", V, want["poly"])
    assert np.max(np.abs(fit_g - fit_w)) <= 1e-6 * np.max(np.abs(fit_w))
    assert np.allclose(got["mse"], want["mse"], rtol=1e-5)


def test_duplicated_columns_minimum_norm(stage, orc):
    """All training stars on one line v = 2 u + 1: the design matrix is rank deficient; A \\ z returns the minimum-norm
    solution, not NaN."""
    rng = np.random.default_rng(77)
    N, r, c, degree = 600, 7, 6, 3
    u = rng.uniform(-1, 1, N)
    v = 2 * u + 1
    pgf = (0.2 + 0.1 * v + 0.05 * v * u)[:, None, None] * (1 + rng.uniform(size=(1, c, r))) + 1e-4 * rng.normal(size=(N, c, r))
    train = rng.permutation(N)[:480]
    got = stage.polyfit(pgf, u, v, train, degree, 2)
    want = orc.polyfit(pgf, u, v, train, degree, 2, nthreads=orc.max_threads())
    assert np.all(np.isfinite(got["poly"]))
    assert got["kept"].tolist() == want["kept"].tolist()
    V = _design(u[train], v[train], degree)
    fit_g, fit_w = np.einsum("nk,kcr->ncr", V, got["poly"]), np.einsum("nk,kcr->ncr", V, want["poly"])
    assert np.max(np.abs(fit_g - fit_w)) <= 1e-6 * np.max(np.abs(fit_w))
    assert coef_close(got["poly"], want["poly"], 1e-6)


def test_median_constraint_with_zero_and_nan_medians(stage, orc):
    """interpolate.jl:118-124 with a per-pixel median of exactly 0 (the first design column vanishes: the rank-revealing
    solve fits WITHOUT a constant term) and with a pixel that is NaN for every star (every coefficient NaN)."""
    pgf, u, v, train, _ = make_case(21, 400, 6, 5, 2, n_out=0)
    pgf[:, 1, 2] -= np.median(pgf[train, 1, 2])                    # odd count below: the median is an element -> exactly 0
    train = train[:301]
    pgf[:, 1, 2] -= np.median(pgf[train, 1, 2])
    assert np.median(pgf[train, 1, 2]) == 0.0
    pgf[:, 3, 3] = 0.0                                              # identically zero pixel
    got = stage.polyfit(pgf, u, v, train, 2, 1, median_constraint=True)
    want = orc.polyfit(pgf, u, v, train, 2, 1, median_constraint=True, nthreads=orc.max_threads())
    assert np.array_equal(got["poly"][0], want["poly"][0])
    assert got["poly"][0, 1, 2] == 0.0 and np.all(got["poly"][:, 3, 3] == 0.0)
    assert coef_close(got["poly"], want["poly"])
    pgf[:, 0, 0] = np.nan
    got = stage.polyfit(pgf, u, v, train, 2, 1, median_constraint=True)
    assert np.all(np.isnan(got["poly"][:, 0, 0])) and np.all(np.isfinite(got["poly"][:, 1, 1]))

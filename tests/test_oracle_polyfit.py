"""CPU: the oracle's polynomial_optimizer / worst_10_percent / polyfit driver / p(u,v) (interpolate.jl:89-147,
shopt.jl:472-577, reader.jl:40-68) against scipy least squares and exact-polynomial known answers."""
import numpy as np
import pytest
import scipy.linalg


def terms(degree):
    return [(p, q) for p in range(degree + 1) for q in range(degree + 1 - p)]     # interpolate.jl:106-114


def design(degree, u, v):
    return np.stack([u ** p * v ** q for p, q in terms(degree)], axis=1)


def np_polyfit(pgf, u, v, train, degree, iterations):
    """shopt.jl:534-577 with numpy lstsq; returns poly (K,c,r), last mse, kept, n_per_round."""
    N, c, r = pgf.shape
    cur = np.array(train)
    npr = []
    for loop in range(iterations):
        npr.append(len(cur))
        A = design(degree, u[cur], v[cur])
        Z = pgf[cur].reshape(len(cur), -1)
        coef = scipy.linalg.lstsq(A, Z, lapack_driver="gelsy")[0]                  # (K, rc)
        mse = ((A @ coef - Z) ** 2).mean(axis=1)
        order = sorted(range(len(cur)), key=lambda i: -mse[i])                      # stable descending
        bad = order[:int(np.ceil(0.10 * len(cur)))]
        if loop < iterations - 1:
            keep = np.ones(len(cur), bool)
            keep[bad] = False
            cur = cur[keep]
    return coef.reshape(-1, c, r), mse, cur, npr


def test_term_order(orc):
    assert orc.num_terms(1) == 3 and orc.num_terms(3) == 10 and orc.num_terms(5) == 21
    # objective_function enumerates (p outer, q inner): coefficient k multiplies u^p v^q
    for deg in (1, 2, 3):
        for k, (p, q) in enumerate(terms(deg)):
            coef = np.zeros(orc.num_terms(deg))
            coef[k] = 1.0
            assert np.isclose(orc.objective_function(coef, 1.7, -0.6, deg), 1.7 ** p * (-0.6) ** q, rtol=1e-15)


@pytest.mark.parametrize("degree", [1, 3, 5])
def test_polynomial_optimizer_matches_lstsq_and_exact(orc, degree):
    rng = np.random.default_rng(degree)
    n = 200
    u, v = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    A = design(degree, u, v)
    truth = rng.normal(size=A.shape[1])
    z = A @ truth
    assert np.allclose(orc.polynomial_optimizer(degree, u, v, z), truth, rtol=1e-9, atol=1e-11)   # exact polynomial
    z = z + 0.1 * rng.normal(size=n)
    want = scipy.linalg.lstsq(A, z, lapack_driver="gelsy")[0]
    assert np.allclose(orc.polynomial_optimizer(degree, u, v, z), want, rtol=1e-9, atol=1e-11)


def test_polynomial_optimizer_overrides(orc):
    rng = np.random.default_rng(11)
    u, v = rng.uniform(-1, 1, 40), rng.uniform(-1, 1, 40)
    z = rng.normal(size=40)
    zn = z.copy()
    zn[[3, 17]] = np.nan
    med = np.median(z[~np.isnan(zn)])
    zr = np.where(np.isnan(zn), med, zn)                                            # interpolate.jl:126-130
    want = scipy.linalg.lstsq(design(2, u, v), zr, lapack_driver="gelsy")[0]
    assert np.allclose(orc.polynomial_optimizer(2, u, v, zn), want, rtol=1e-9)
    c = orc.polynomial_optimizer(2, u, v, z, median_constraint=True)                # :118-124, 139-141
    assert c[0] == np.median(z)
    c4 = orc.polynomial_optimizer(1, u[:4], v[:4], z[:4])                           # :142-145
    assert c4[0] == np.median(z[:4]) and np.all(c4[1:] == 0)


def test_median(orc):
    assert orc.median(np.array([3.0, np.nan, 1.0, 2.0])) == 2.0
    assert orc.median(np.array([4.0, 1.0, 3.0, 2.0])) == 2.5


def test_worst_10_percent(orc):
    err = np.array([1.0, 5.0, 5.0, 0.5, 5.0, 2.0, 0.1, 0.2, 0.3, 0.4, 0.6])       # n=11 -> ceil(1.1)=2
    assert orc.worst_10_percent(err).tolist() == [1, 2]                           # ties keep index order (stable sort)
    assert len(orc.worst_10_percent(np.arange(100.0))) == 10
    assert orc.worst_10_percent(np.arange(100.0)).tolist() == list(range(99, 89, -1))
    e = np.ones(50)
    e[7] = 100.0
    assert orc.worst_10_percent(e, sigma_clip=True).tolist() == [7]               # mean + 3 sigma (shopt.jl:522-532)


@pytest.mark.parametrize("degree,iterations", [(1, 3), (3, 3), (2, 1)])
def test_polyfit_driver(orc, degree, iterations):
    rng = np.random.default_rng(degree * 10 + iterations)
    N, r, c = 120, 6, 5
    u, v = rng.uniform(-1, 1, N), rng.uniform(-1, 1, N)
    K = orc.num_terms(degree)
    truth = rng.normal(size=(K, c, r))
    pgf = np.einsum("nk,kcr->ncr", design(degree, u, v), truth) + 0.01 * rng.normal(size=(N, c, r))
    pgf[rng.choice(N, 6, replace=False)] += 1.0                                    # outlier stars
    train = rng.permutation(N)[:100]
    got = orc.polyfit(pgf, u, v, train, degree, iterations)
    poly, mse, kept, npr = np_polyfit(pgf, u, v, train, degree, iterations)
    assert got["n_per_round"].tolist() == npr
    assert got["kept"].tolist() == kept.tolist()                                   # selection bit-exact
    assert np.allclose(got["poly"], poly, rtol=1e-8, atol=1e-10)
    assert np.allclose(got["mse"], mse, rtol=1e-8)


def test_psf_at(orc):
    rng = np.random.default_rng(2)
    poly = rng.normal(size=(10, 4, 5)) + 3.0
    star = orc.psf_at(poly, 3, 0.3, -0.7)
    want = np.einsum("k,kcr->cr", design(3, np.array([0.3]), np.array([-0.7]))[0], poly)
    assert np.allclose(star, want / want.sum(), rtol=1e-12)                        # reader.jl:40-68


def test_field_interpolation_is_the_least_squares_solution(stage):
    """shopt.jl:422-466 minimises interpCostS (interpolate.jl:7-19) by LBFGS from rand(10); the cost is a convex
    quadratic, so the product's direct QR solution, numpy's SVD-based lstsq and a quasi-Newton run from a random
    start must all land on the same coefficients.  Host code: runs without a device."""
    from scipy.optimize import minimize
    from oracle import stage_oracle
    rng = np.random.default_rng(12)
    n = 500
    u, v = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    s = 0.30 + 0.05 * u - 0.03 * v + 0.02 * u * v + 0.01 * u ** 3 + 1e-3 * rng.normal(size=n)
    g1 = 0.08 + 0.03 * u - 0.02 * v * v + 1e-3 * rng.normal(size=n)
    g2 = -0.03 + 0.03 * v + 0.02 * u * u + 1e-3 * rng.normal(size=n)
    coef, rank = stage.field_interp(u, v, s, g1, g2)
    want = stage_oracle.field_interp(u, v, s, g1, g2)
    assert rank == 10 and np.allclose(coef, want, rtol=1e-9, atol=1e-12)

    def cost(p):                                   # interpolate.jl:8-18 as written
        I = p[0]*u**3 + p[1]*v**3 + p[2]*u**2*v + p[3]*v**2*u + p[4]*u**2 + p[5]*v**2 + p[6]*u*v + p[7]*u + p[8]*v + p[9]
        return np.sum((I - s) ** 2)
    res = minimize(cost, rng.uniform(size=10), method="L-BFGS-B", options=dict(gtol=1e-12, ftol=1e-30, maxiter=5000))
    assert np.allclose(res.x, coef[0], atol=2e-5)
    assert np.allclose(stage.field_eval(coef[0], u[:7], v[:7]),
                       [np.dot(want[0], [a**3, b**3, a*a*b, b*b*a, a*a, b*b, a*b, a, b, 1.0])
                        for a, b in zip(u[:7], v[:7])], rtol=1e-12)
    # rank-deficient input (all stars on one line): the dependent coefficients are zeroed, the fit is still exact
    coef, rank = stage.field_interp(u, u, 2.0 + 3.0 * u)
    assert rank < 10 and np.allclose(stage.field_eval(coef[0], u, u), 2.0 + 3.0 * u, atol=1e-9)

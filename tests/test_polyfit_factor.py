"""CPU: the K x K part of the product's  A \\ z  (shopt_polyfit_factor_host -- the same polyfit_factor.h code the device
kernel runs) against LAPACK's column-pivoted QR with the reference's rank rule and minimum-norm completion
(interpolate.jl:138 through Julia's `\\`), on well-conditioned, raw-sky-coordinate, underdetermined and rank-deficient
designs.  No device needed."""
import ctypes as C

import numpy as np
import pytest
import scipy.linalg as sl

from shopt_b200 import _lib


def terms(deg):
    return [(p, q) for p in range(deg + 1) for q in range(deg + 1 - p)]


def design(u, v, deg):
    return np.stack([u ** p * v ** q for p, q in terms(deg)], axis=1)


def lapack_solve(A, z):
    """qr(A, ColumnNorm()) \\ z: rank = #{k: |R_kk| > min(n,K) eps |R_11|}, minimum-norm solution when rank < K."""
    n, K = A.shape
    if n == K:
        return np.linalg.solve(A, z), K
    Q, R, P = sl.qr(A, mode="economic", pivoting=True)
    mn = min(n, K)
    d = np.abs(np.diag(R))
    rank = 1
    while rank < mn and d[rank] > mn * np.finfo(float).eps * d[0]:
        rank += 1
    c = Q.T @ z
    if rank == K:
        y = sl.solve_triangular(R, c)
    else:
        R1 = R[:rank, :]
        y = R1.T @ np.linalg.solve(R1 @ R1.T, c[:rank])
    x = np.zeros(K)
    x[P] = y
    return x, rank


def product_solve(u, v, Z, deg):
    """Normal equations in the centred basis (what normal_kernel accumulates) -> shopt_polyfit_factor_host."""
    n = len(u)
    u0, v0 = u.mean(), v.mean()
    su = np.sqrt(3 * u.var()) if u.var() > 1e-12 * u0 * u0 and u.var() > 0 else 1.0
    sv = np.sqrt(3 * v.var()) if v.var() > 1e-12 * v0 * v0 and v.var() > 0 else 1.0
    ctr = np.array([u0, 1 / su, v0, 1 / sv])
    Vc = design((u - u0) / su, (v - v0) / sv, deg)
    G, b = np.ascontiguousarray(Vc.T @ Vc), np.ascontiguousarray(Vc.T @ Z)
    coef, info = np.zeros_like(b), C.c_int32(0)
    rank = _lib.lib().shopt_polyfit_factor_host(G.ctypes.data_as(_lib.dp), deg, ctr.ctypes.data_as(_lib.dp), n,
                                                b.ctypes.data_as(_lib.dp), b.shape[1], coef.ctypes.data_as(_lib.dp), C.byref(info))
    return coef, rank, info.value


@pytest.mark.parametrize("deg,uc,vc,w,n", [
    (1, 150.0, 2.0, 0.1, 2000), (3, 150.0, 2.0, 0.1, 2000), (3, 0.0, 0.0, 1.0, 2000), (5, 0.0, 0.0, 1.0, 2000),
    (3, 150.1, 2.2, 0.1, 500), (2, 150.0, 2.0, 0.1, 100), (5, 150.0, 2.0, 0.1, 3000), (3, 0.0, 0.0, 1.0, 7),
    (2, 0.0, 0.0, 1.0, 6), (8, 0.0, 0.0, 1.0, 4000), (3, -35.0, 60.0, 0.5, 800)])
def test_matches_lapack_pivoted_qr(deg, uc, vc, w, n):
    rng = np.random.default_rng(deg * 100 + n)
    u, v = uc + w * rng.uniform(-1, 1, n), vc + w * rng.uniform(-1, 1, n)
    a, b = (u - uc) / w, (v - vc) / w
    Z = np.stack([0.3 + 0.05 * a - 0.03 * b + 0.02 * a ** 3 + 1e-5 * rng.normal(size=n),
                  1e-3 * (1 + a * b) + 1e-6 * rng.normal(size=n)], axis=1)
    A = design(u, v, deg)
    coef, rank, info = product_solve(u, v, Z, deg)
    for j in range(Z.shape[1]):
        want, wrank = lapack_solve(A, Z[:, j])
        assert rank == wrank
        fit_w, fit_g = A @ want, A @ coef[:, j]
        assert np.max(np.abs(fit_w - fit_g)) <= 1e-6 * np.max(np.abs(fit_w))          # fitted values: the bar
        if np.linalg.cond(A) < 1e6:
            assert np.allclose(coef[:, j], want, rtol=1e-8, atol=1e-9 * np.abs(want).max())
    K = A.shape[1]
    assert bool(info & _lib.POLY_RANK_DEFICIENT) == (rank < K)


def test_exactly_dependent_columns_give_the_minimum_norm_solution():
    rng = np.random.default_rng(3)
    n = 300
    u, v = np.full(n, 0.5), rng.uniform(-1, 1, n)              # every star at the same u: u^p columns are multiples of 1
    Z = (0.2 + 0.1 * v + 0.05 * v * v)[:, None]
    coef, rank, info = product_solve(u, v, Z, 2)
    want, wrank = lapack_solve(design(u, v, 2), Z[:, 0])
    assert rank == wrank == 3 and info & _lib.POLY_RANK_DEFICIENT
    assert np.allclose(coef[:, 0], want, atol=1e-12)
    u = rng.uniform(-1, 1, n); v = 2 * u + 1                    # stars on a line
    Z = (0.2 + 0.1 * v + 0.05 * v * u)[:, None]
    A = design(u, v, 3)
    coef, rank, info = product_solve(u, v, Z, 3)
    want, wrank = lapack_solve(A, Z[:, 0])
    assert rank == wrank == 4
    assert np.allclose(A @ coef[:, 0], A @ want, atol=1e-12) and np.allclose(coef[:, 0], want, atol=1e-10)


def test_non_finite_normal_equations_are_reported():
    G = np.eye(3); G[1, 1] = np.nan
    b, coef, info = np.ones((3, 1)), np.zeros((3, 1)), C.c_int32(0)
    rank = _lib.lib().shopt_polyfit_factor_host(G.ctypes.data_as(_lib.dp), 1, None, 10, b.ctypes.data_as(_lib.dp), 1,
                                                coef.ctypes.data_as(_lib.dp), C.byref(info))
    assert rank == 0 and info.value & _lib.POLY_NONFINITE and np.all(np.isnan(coef))

"""Generates tests/golden/*.npz.  The reference ships no golden vectors and cannot run here (Julia absent), so these
come from INDEPENDENT numpy/scipy restatements of the reference lines cited below -- not from oracle/ and not from the
CUDA path -- and pin both.  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np
import scipy.linalg
import scipy.optimize

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from shopt_b200 import synth  # noqa: E402  (input generator only)


def e_to_g(e1, e2):                                   # analyticLBFGS.jl:10-18
    ell = np.sqrt(e1 ** 2 + e2 ** 2)
    normG = np.sqrt(1 + 0.5 * (1 / ell ** 2 - np.sqrt(4 / ell ** 2 + 1 / ell ** 4)))
    return e1 * normG / ell, e2 * normG / ell


def cost(stamp, n_a, x):                              # analyticLBFGS.jl:8-46, radialProfiles.jl:5-9
    c, r = stamp.shape
    g1, g2 = e_to_g(x[1], x[2])
    r0, c0 = (r - n_a) // 2, (c - n_a) // 2
    crop = stamp[c0:c0 + n_a, r0:r0 + n_a]
    u = np.arange(1, n_a + 1) - n_a / 2.0
    du, dv = u[None, :], u[:, None]
    w1 = (1 + g1) * du + g2 * dv
    w2 = g2 * du + (1 - g1) * dv
    f = np.exp(-((x[0] ** 2) ** 2 / (1 - g1 * g1 - g2 * g2)) * (w1 * w1 + w2 * w2))
    res = f / f.sum() - crop
    return 0.5 * np.nansum(res * res)


def num_grad(stamp, n_a, x, h=1e-6):
    g = np.zeros(3)
    for j in range(3):
        xp, xm = x.copy(), x.copy()
        xp[j] += h
        xm[j] -= h
        g[j] = (cost(stamp, n_a, xp) - cost(stamp, n_a, xm)) / (2 * h)
    return g


def pca_image(img, ncomp, pratio=0.99):               # pca.jl:1-18 via LAPACK
    X = img.T
    mu = X.mean(axis=1, keepdims=True)
    Z = X - mu
    U, S, _ = np.linalg.svd(Z, full_matrices=False)
    var = S ** 2 / (X.shape[1] - 1)
    k, acc = 1, var[0]
    while k < len(var) and k < ncomp and acc < pratio * var.sum():
        acc += var[k]
        k += 1
    return (U[:, :k] @ (U[:, :k].T @ Z) + mu).T, k


def design(degree, u, v):                             # interpolate.jl:106-114, 131-135
    return np.stack([u ** p * v ** q for p in range(degree + 1) for q in range(degree + 1 - p)], axis=1)


def main():
    # ---- analytic: cost, finite-difference gradient, minimiser (scipy BFGS polished by Nelder-Mead) ----
    N, r, n_a = 6, 27, 24
    cat = synth.make_catalog(N, r, n_a, seed=7)
    st, x0 = cat["stamps"].numpy(), cat["x0_model"].numpy()
    st[1, 5, 6] = np.nan
    costs = np.array([cost(st[i], n_a, x0[i]) for i in range(N)])
    grads = np.array([num_grad(st[i], n_a, x0[i]) for i in range(N)])
    mins = []
    for i in range(N):
        res = scipy.optimize.minimize(lambda x: cost(st[i], n_a, x), np.array([0.55, 0.1, 0.05]), method="BFGS",
                                      options=dict(gtol=1e-12))
        res = scipy.optimize.minimize(lambda x: cost(st[i], n_a, x), res.x, method="Nelder-Mead",
                                      options=dict(xatol=1e-11, fatol=1e-20, maxiter=5000))
        g1, g2 = e_to_g(res.x[1], res.x[2])
        mins.append([res.x[0] ** 2, g1, g2])
    np.savez_compressed(os.path.join(HERE, "analytic.npz"), stamps=st, x0=x0, n_a=n_a, cost=costs, grad_fd=grads,
                        minimum=np.array(mins))
    # ---- PCA reconstructions ----
    cat = synth.make_catalog(4, 33, 32, seed=11, noise=1e-4)
    img = cat["stamps"].numpy()
    rng = np.random.default_rng(5)
    img = np.concatenate([img, (rng.normal(size=(1, 33, 33)))])
    rec = [pca_image(np.where(np.isnan(x), 0.001, x), 50) for x in img]
    np.savez_compressed(os.path.join(HERE, "pca.npz"), stamps=img, recon=np.stack([x[0] for x in rec]),
                        k=np.array([x[1] for x in rec]))
    # ---- polyfit rounds ----
    N, rr, cc, degree = 150, 5, 4, 3
    u, v = rng.uniform(-1, 1, N), rng.uniform(-1, 1, N)
    truth = rng.normal(size=(10, cc, rr))
    pgf = np.einsum("nk,kcr->ncr", design(degree, u, v), truth) + 0.02 * rng.normal(size=(N, cc, rr))
    pgf[rng.choice(N, 8, replace=False)] += 1.5
    train = rng.permutation(N)[:120]
    cur, sizes = train.copy(), []
    for loop in range(3):
        sizes.append(len(cur))
        A = design(degree, u[cur], v[cur])
        Zm = pgf[cur].reshape(len(cur), -1)
        coef = scipy.linalg.lstsq(A, Zm, lapack_driver="gelsy")[0]
        mse = ((A @ coef - Zm) ** 2).mean(axis=1)
        if loop < 2:
            order = sorted(range(len(cur)), key=lambda i: -mse[i])
            keep = np.ones(len(cur), bool)
            keep[order[:int(np.ceil(0.1 * len(cur)))]] = False
            cur = cur[keep]
    np.savez_compressed(os.path.join(HERE, "polyfit.npz"), pgf=pgf, u=u, v=v, train=train, degree=degree,
                        poly=coef.reshape(10, cc, rr), mse=mse, kept=cur, sizes=np.array(sizes))
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()

"""numpy mirror of csrc/pca_eig.cuh (same recurrences) to check the maths before spending GPU time"""
import numpy as np, sys
rng = np.random.default_rng(0)
EPS = np.finfo(float).eps

def tridiag(A):
    n = A.shape[0]; A = A.copy(); d = np.zeros(n); e = np.zeros(n); tau = np.zeros(n); H = np.zeros((n, n))
    for k in range(n - 2):
        m0 = k + 1
        x = A[:, k]
        ss = np.sum(x[m0 + 1:] ** 2); x0 = x[m0]
        alpha, beta, u0 = x0, 0.0, 0.0
        if ss > 2.2e-308:
            nrm = np.sqrt(x0 * x0 + ss); alpha = -nrm if x0 > 0 else nrm; u0 = x0 - alpha; beta = 2.0 / (u0 * u0 + ss)
        u = np.zeros(n)
        if beta != 0: u[m0] = u0; u[m0 + 1:] = x[m0 + 1:]
        d[k] = x[k]; e[k] = alpha; tau[k] = beta
        if beta != 0:
            p = beta * (A[:, m0:] @ u[m0:]); p[:m0] = 0
            K = 0.5 * beta * (u @ p); q = p - K * u; q[:m0] = 0
            A[m0:, m0:] -= np.outer(u[m0:], q[m0:]) + np.outer(q[m0:], u[m0:])
            H[m0:, k] = u[m0:]
    d[n - 2] = A[n - 2, n - 2]; e[n - 2] = A[n - 1, n - 2]; d[n - 1] = A[n - 1, n - 1]
    return d, e, tau, H

def sturm(d, e2, x, pivmin):
    q = d[0] - x
    if abs(q) < pivmin: q = -pivmin
    cnt = int(q < 0)
    for i in range(1, len(d)):
        q = d[i] - x - e2[i - 1] / q
        if abs(q) < pivmin: q = -pivmin
        cnt += q < 0
    return cnt

def top_eigs(d, e, m):
    n = len(d); e2 = e * e
    rad = np.abs(e) + np.concatenate([[0], np.abs(e[:-1])])
    lo0, hi0 = (d - rad).min(), (d + rad).max(); tn = max(abs(lo0), abs(hi0))
    lo0 -= 2 * EPS * n * tn; hi0 += 2 * EPS * n * tn
    pivmin = 2.2e-308 * max(1.0, tn * tn)
    lam = np.zeros(m)
    for j in range(m):
        want = n - j; lo, hi = lo0, hi0
        for it in range(26):
            w = (hi - lo) / 9.0
            b = [sturm(d, e2, lo + w * (s + 1), pivmin) >= want for s in range(8)]
            s = b.index(True) if any(b) else 8
            nlo = lo if s == 0 else lo + w * s; nhi = hi if s == 8 else lo + w * (s + 1)
            lo, hi = nlo, nhi
            if hi - lo <= 2 * EPS * max(abs(lo), abs(hi)) + 2 * pivmin: break
        lam[j] = 0.5 * (lo + hi)
    return lam, tn

def invit(d, e, lamj, j, tnorm):
    n = len(d); tiny = max(EPS * tnorm, 2.2e-308)
    F = np.zeros(n); RD = np.zeros(n); SW = np.zeros(n, int)
    di = d[0] - lamj; dui = e[0]
    for i in range(n - 1):
        dli, dnext, dunext = e[i], d[i + 1] - lamj, e[i + 1]
        if abs(di) >= abs(dli):
            if di == 0: di = tiny
            f = dli / di; F[i] = f; RD[i] = 1 / di; SW[i] = 0
            di = dnext - f * dui; dui = dunext
        else:
            f = di / dli; F[i] = f; RD[i] = 1 / dli; SW[i] = 1
            di, dui = dui - f * dnext, -f * dunext
    if abs(di) < tiny: di = -tiny if di < 0 else tiny
    RD[n - 1] = 1 / di
    v = 0.5 + rng.random(n)
    scale = 1.0
    for it in range(3):
        if it > 0:
            bi = v[0] * scale
            for i in range(n - 1):
                bn = v[i + 1] * scale; f = F[i]
                if SW[i]: v[i] = bn; bi = bi - f * bn
                else: v[i] = bi; bi = bn - f * bi
            v[n - 1] = bi
        x1 = v[n - 1] * RD[n - 1]; x2 = 0.0; mx = abs(x1); v[n - 1] = x1
        for i in range(n - 2, -1, -1):
            if SW[i]: du = d[i + 1] - lamj; du2 = e[i + 1]
            else: du = (-F[i - 1] * e[i]) if (i > 0 and SW[i - 1]) else e[i]; du2 = 0.0
            x = (v[i] - du * x1 - du2 * x2) * RD[i]; v[i] = x; mx = max(mx, abs(x)); x2 = x1; x1 = x
        scale = 1 / mx
    v = v * scale; v /= np.linalg.norm(v)
    t = (d - lamj) * v; t[1:] += e[:-1] * v[:-1]; t[:-1] += e[:-1] * v[1:]
    return v, np.linalg.norm(t)

def run(Z, ncomp=50, pratio=0.99):
    r, c = Z.shape; C = Z @ Z.T
    d, e, tau, H = tridiag(C)
    m = min(ncomp, r, c)
    lam, tn = top_eigs(d, e, m)
    thres = d.sum() / (c - 1) * pratio; kk = 1; acc = lam[0] / (c - 1)
    while kk < m and acc < thres: acc += lam[kk] / (c - 1); kk += 1
    V = np.zeros((r, kk)); resmax = 0
    for j in range(kk):
        V[:, j], res = invit(d, e, lam[j], j, tn); resmax = max(resmax, res / tn)
    G = V.T @ V - np.eye(kk)
    for a in range(kk):
        for k in range(r - 3, -1, -1):
            if tau[k] != 0: V[:, a] -= tau[k] * (H[:, k] @ V[:, a]) * H[:, k]
    return lam, kk, V, resmax, np.abs(G).max()

for trial in range(3):
    r = c = 41 + 10 * trial
    yy, xx = np.mgrid[:r, :c]; g = np.exp(-((yy - r / 2) ** 2 + (xx - c / 2) ** 2) / 30.0); g /= g.sum()
    X = g + 1e-5 * rng.standard_normal((r, c)); X[rng.random((r, c)) < 0.01] = 0.001
    Z = X - X.mean(axis=1, keepdims=True)
    lam, k, U, res, orth = run(Z)
    Uref, s, _ = np.linalg.svd(Z); lref = s ** 2
    P, Pref = U @ U.T, Uref[:, :k] @ Uref[:, :k].T
    print("n=%d k=%d  eig rel err %.2e  T-res %.2e  orth %.2e  projector err %.2e" % (r, k, np.abs(lam[:k] - lref[:k]).max() / lref[0], res, orth, np.abs(P - Pref).max()))
print("flat spectra")
for trial in range(4):
    r = c = 48 + 16 * trial
    X = rng.standard_normal((r, c)) * (1.0 / (1 + np.arange(c)))[None, :] ** (0.3 * trial) * 1e-4
    if trial == 3: X[:, 5] = X[:, 6]      # exactly repeated column
    Z = X - X.mean(axis=1, keepdims=True)
    lam, k, U, res, orth = run(Z, ncomp=50)
    Uref, s, _ = np.linalg.svd(Z); lref = s ** 2
    P, Pref = U @ U.T, Uref[:, :k] @ Uref[:, :k].T
    gap = (lref[k - 1] - lref[k]) / lref[0]
    print("n=%d k=%d  eig rel err %.2e  T-res %.2e  orth %.2e  projector err %.2e (boundary gap %.1e)" % (r, k, np.abs(lam[:k] - lref[:k]).max() / lref[0], res, orth, np.abs(P - Pref).max(), gap))

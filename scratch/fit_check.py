import ctypes as C, numpy as np, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from shopt_b200 import synth
from oracle import shopt_oracle as orc
L = C.CDLL("shopt_b200/libshopt_b200.so")
dp = C.POINTER(C.c_double); ip = C.POINTER(C.c_int32)
L.shopt_fit_analytic_batch.restype = C.c_int
L.shopt_fit_analytic_batch.argtypes = [dp, C.c_int, C.c_int, C.c_int64, C.c_int, dp, C.c_double, C.c_int, dp, dp, dp, ip, ip]
L.shopt_last_error.restype = C.c_char_p
L.shopt_last_timing.argtypes=[dp]
def run(N, r, n_a, **kw):
    cat = synth.make_catalog(N, r, n_a, **kw)
    st = cat["stamps"].numpy(); x0 = cat["x0_model"].numpy()
    s = np.zeros(N); g1 = np.zeros(N); g2 = np.zeros(N); status = np.zeros(N, np.int32); iters = np.zeros(N, np.int32)
    p = lambda a, t=dp: a.ctypes.data_as(t)
    t0 = time.time()
    rc = L.shopt_fit_analytic_batch(p(st), r, r, N, n_a, p(x0), 1e-6, 1000, p(s), p(g1), p(g2), p(status, ip), p(iters, ip))
    t1 = time.time()
    ms = (C.c_double*3)(); L.shopt_last_timing(ms)
    print("rc", rc, L.shopt_last_error(), "wall %.3fs" % (t1 - t0), "ms", list(ms))
    n_o = min(N, 256)
    t0 = time.time()
    os_, og1, og2, ostat, oit = orc.fit_analytic_batch(st[:n_o], n_a, x0[:n_o], nthreads=orc.max_threads())
    t1 = time.time()
    print("oracle %d stars %.3fs threads %d" % (n_o, t1 - t0, orc.max_threads()))
    rel = lambda a, b: np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))
    print("N", N, "r", r, "n_a", n_a, "relerr s,g1,g2:", rel(s[:n_o], os_), rel(g1[:n_o], og1), rel(g2[:n_o], og2))
    print("status eq", np.array_equal(status[:n_o], ostat), "iters eq", np.array_equal(iters[:n_o], oit), "iters mean", iters.mean(), "status hist", np.bincount(status + 1))
    bad = np.nonzero(np.abs(s[:n_o]-os_) > 1e-6*np.abs(os_))[0]
    print("n_bad", len(bad), bad[:10], [(s[i], os_[i], iters[i], oit[i]) for i in bad[:5]])
    print("truth err", np.abs(s - cat["truth"][:, 0].numpy()).max())
    print("kernel stars/s", N / (ms[1] * 1e-3))
for (N,r,n_a) in [(300,75,75),(40,75,75),(300,77,75),(300,75,73),(300,76,76),(300,80,70),(300,78,78),(300,67,67)]:
    run(N,r,n_a)

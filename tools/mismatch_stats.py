import numpy as np, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import synth, stage
from oracle import shopt_oracle as orc
for (N, r, n_a) in [(2000, 75, 75), (2000, 76, 76), (2000, 131, 64), (1000, 140, 128), (2000, 41, 32)]:
    cat = synth.make_catalog(N, r, n_a)
    st = cat["stamps"].numpy(); x0 = cat["x0_model"].numpy()
    got = stage.fit_analytic_batch(st, n_a, x0)
    want = orc.fit_analytic_batch(st, n_a, x0, nthreads=16)
    bad = np.nonzero((np.abs(got[0] - want[0]) > 1e-6 * np.abs(want[0])) | (got[4] != want[4]))[0]
    f, g = stage.analytic_cost_grad_batch(st, n_a, x0)
    fo = np.array([orc.cost(st[i], n_a, x0[i]) for i in range(200)]); go = np.array([orc.grad(st[i], n_a, x0[i]) for i in range(200)])
    print(N, r, n_a, "mismatch", len(bad), "sigma0 of mismatches", np.round(np.sort(x0[bad, 0]), 2).tolist()[:30],
          "| f rel %.1e g rel %.1e" % (np.max(np.abs(f[:200] - fo) / fo), np.max(np.abs(g[:200] - go) / np.abs(go).max(axis=1, keepdims=True))))

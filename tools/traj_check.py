import numpy as np, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import synth, stage
from oracle import shopt_oracle as orc
np.set_printoptions(precision=17)
for (N, r, n_a, stars) in [(40, 75, 75, [8, 23]), (300, 75, 73, [92, 31])]:
    cat = synth.make_catalog(N, r, n_a)
    st = cat["stamps"].numpy(); x0 = cat["x0_model"].numpy()
    for i in stars:
        print("== star", i, "x0", x0[i])
        for k in range(0, 14):
            s, g1, g2, status, iters = stage.fit_analytic_batch(st[i:i+1], n_a, x0[i:i+1], 1e-6, k)
            o = orc.lbfgs_fit(st[i], n_a, x0[i], 1e-6, k)
            print(k, "gpu", s[0], g1[0], status[0], iters[0], "| orc", o["p"][0], o["p"][1], o["status"], o["iterations"], o["evals"], o["trials"],
                  "rel %.2e" % (abs(s[0]-o["p"][0])/abs(o["p"][0])))

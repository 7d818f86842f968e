import numpy as np, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import synth, stage
from oracle import shopt_oracle as orc
for (N, r, n_a) in [(64, 75, 75), (64, 76, 76), (64, 77, 75), (64, 77, 70), (64, 80, 75), (3, 75, 75), (64, 131, 64)]:
    cat = synth.make_catalog(N, r, n_a)
    st = cat["stamps"].numpy(); x = cat["x0_model"].numpy()
    f, g = stage.analytic_cost_grad_batch(st, n_a, x)
    fo = np.array([orc.cost(st[i], n_a, x[i]) for i in range(N)])
    go = np.array([orc.grad(st[i], n_a, x[i]) for i in range(N)])
    rf = np.abs(f - fo) / np.abs(fo)
    rg = np.abs(g - go) / np.abs(go).max(axis=1, keepdims=True)
    print(N, r, n_a, "f rel max %.3e  g rel max %.3e" % (rf.max(), rg.max()), "worst star", rf.argmax(), rg.max(axis=1).argmax())

import numpy as np, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import synth, pipeline
from oracle import stage_oracle, shopt_oracle as orc
N, r, n_a, degree = 160, 41, 32, 1
cat = synth.make_catalog(N, r, n_a)
a = {k: cat[k].numpy() for k in ("stamps", "u", "v", "x0_model", "x0_learned", "perm")}
p = pipeline.default_params(n_a=n_a, degree=degree, iterations=3, training_ratio=0.8)
got = pipeline.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"])
want = stage_oracle.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"], nthreads=16)
for nm in ("model", "learned"):
    d = np.abs(got[nm] - want[nm]) / np.abs(want[nm])
    for row in range(3):
        i = d[row].argmax()
        print(nm, row, "worst rel %.2e at star %d: %r vs %r" % (d[row, i], i, got[nm][row, i], want[nm][row, i]), "k", want["pca_k"][i])
pd_ = np.abs(got["pgf"] - want["pgf"]).max(axis=(1, 2)) / np.abs(want["pgf"]).max(axis=(1, 2))
print("pgf worst rel-to-peak %.2e" % pd_.max(), "k hist", np.bincount(want["pca_k"]))
sc = np.abs(want["poly"]).max(axis=(1, 2), keepdims=True)
print("poly worst %.2e" % (np.abs(got["poly"] - want["poly"]) / sc).max())

import numpy as np, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import synth, stage
from oracle import shopt_oracle as orc
for (N, r, n_a) in [(160, 41, 32), (2000, 41, 32), (2000, 33, 32), (2000, 70, 64), (2000, 75, 75), (1000, 140, 128)]:
    cat = synth.make_catalog(N, r, n_a, device="cuda")
    base = stage.fit_analytic_batch(cat["stamps"], n_a, cat["x0_model"])
    torch.cuda.synchronize()
    ndiff = 0
    for rep in range(4):
        perm = torch.randperm(N, device="cuda")
        got = stage.fit_analytic_batch(cat["stamps"][perm].contiguous(), n_a, cat["x0_model"][perm].contiguous())
        torch.cuda.synchronize()
        ndiff += int((base[0][perm] != got[0]).sum())
    n_o = min(N, 300)
    want = orc.fit_analytic_batch(cat["stamps"][:n_o].cpu().numpy(), n_a, cat["x0_model"][:n_o].cpu().numpy(), nthreads=16)
    s = base[0][:n_o].cpu().numpy()
    bad = np.nonzero(np.abs(s - want[0]) > 1e-6 * np.abs(want[0]))[0]
    print(N, r, n_a, "nondeterministic stars:", ndiff, "| vs oracle bad:", bad.tolist(), [float(cat["x0_model"][i, 0]) for i in bad])

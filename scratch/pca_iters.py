"""Histogram of the subspace iterations the PCA fast path used (SHOPT_PCA_DEBUG bit 0) and timing, on the bench
catalogue and on the masked (config-4 style) catalogue."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from shopt_b200 import stage, synth
for kw in ({}, dict(frac_masked=0.01), dict(frac_masked=0.01, frac_contaminated=0.02, frac_bad_s=0.01)):
    cat = synth.make_catalog(8000, 131, 64, **kw)
    x = cat["stamps"].cuda()
    stage.pca_batch(x[:500], 50, 0.99, 3, False)
    torch.cuda.synchronize(); t = time.time()
    pgf, k, status = stage.pca_batch(x, 50, 0.99, 3, False)
    torch.cuda.synchronize(); dt = time.time() - t
    st = status.cpu().numpy()
    print(kw, "ms per 100k stars: %.1f" % (dt * 1e3 * 100000 / 8000))
    print("  iterations at acceptance (0 = fell back to Jacobi):", np.bincount(st >> 8))
    print("  k:", np.bincount(k.cpu().numpy()))

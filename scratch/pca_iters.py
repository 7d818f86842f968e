"""Histogram of the subspace iterations the PCA fast path used (SHOPT_PCA_DEBUG=1) on the bench catalogue."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from shopt_b200 import stage, synth
cat = synth.make_catalog(4000, 131, 64)
pgf, k, status = stage.pca_batch(cat["stamps"].cuda(), 50, 0.99, 3, False)
st = status.cpu().numpy()
print("iterations at acceptance (0 = fell back to Jacobi):", np.bincount(st >> 8))
print("k:", np.bincount(k.cpu().numpy()))

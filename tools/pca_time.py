"""pca_kernel on a masked (config-4 style) catalogue: CUDA-event timing (SHOPT_PCA_DEBUG=16 adds cycles per phase)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from shopt_b200 import stage, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
masked = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
cat = synth.make_catalog(n, 131, 64, frac_masked=masked)
x = cat["stamps"].cuda()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(2):
    torch.cuda.synchronize(); e0.record()
    pgf, k, status = stage.pca_batch(x, 50, 0.99, 3, False)
    e1.record(); torch.cuda.synchronize()
    del pgf
print("masked %.3f: ms per 100k stars: %.1f" % (masked, e0.elapsed_time(e1) * 100000 / n))

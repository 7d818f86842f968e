import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import synth, stage
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
cat = synth.make_catalog(N, 131, 64, device="cuda")
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = stage.fit_analytic_batch(cat["stamps"], 64, cat["x0_model"])
    e1.record(); torch.cuda.synchronize()
    print("fit %d stars: %.2f ms, iters mean %.2f" % (N, e0.elapsed_time(e1), out[4].float().mean().item()))

#!/usr/bin/env python3
"""One small stage per polyfit tiling under compute-sanitizer (memcheck / racecheck / synccheck):

    compute-sanitizer --tool racecheck python tools/sanitize_stage.py

Covers the cp.async ring kernels (normal_ring_kernel, mse_ring_kernel at 4 pixels per thread for K = 6 and K = 21,
1 pixel for K = 28), the downdated rounds, the PCA column ring (SHOPT_PCA_RING=1 forces it on small stamps) and the
compaction that lists the rejected stars.  Prints one line per case; the sanitizer's summary follows on exit."""
import os
import sys

os.environ.setdefault("SHOPT_PCA_RING", "1")
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

from shopt_b200 import pipeline, synth  # noqa: E402


def main():
    for (n, r, n_a, degree) in ((96, 33, 32, 2), (80, 41, 32, 5), (72, 33, 32, 6)):
        cat = synth.make_catalog(n, r, n_a, frac_contaminated=0.05)
        p = pipeline.default_params(n_a=n_a, degree=degree, iterations=3, training_ratio=0.9)
        out = pipeline.run_stage(p, *(cat[k].numpy() for k in ("stamps", "u", "v", "x0_model", "x0_learned", "perm")))
        print("stage n=%d r=%d degree=%d: rounds %s, poly finite %s" % (
            n, r, degree, out["n_per_round"].tolist(), bool((out["poly"] == out["poly"]).all())), flush=True)


if __name__ == "__main__":
    main()

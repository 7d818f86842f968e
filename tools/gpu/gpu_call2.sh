#!/bin/bash
# masked-PCA work: parity of the pixel-grid tests, then per-phase cycles of the masked catalogue
set -u
O=gpurun_out; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_pixelgrid.py tests/test_gpu_baseline_shapes.py -x -q ) > $O/t2_tests.log 2>&1
echo "tests rc=$?"; tail -5 $O/t2_tests.log
SHOPT_PCA_DEBUG=16 timeout 300 python tools/pca_time.py 14800 0.01 > $O/t2_pca_masked.log 2>&1; tail -2 $O/t2_pca_masked.log
timeout 300 python tools/pca_time.py 29600 0.0 > $O/t2_pca_unmasked.log 2>&1; cat $O/t2_pca_unmasked.log

#!/bin/bash
# two-stage PCA (pca_power_kernel + pca_kernel on the left-over list): parity, then timings
set -u
O=gpurun_out; mkdir -p $O
( timeout 1200 python -m pytest tests/test_gpu_pixelgrid.py tests/test_gpu_stage.py tests/test_gpu_baseline_shapes.py tests/test_gpu_modes.py tests/test_gpu_ingest.py -x -q ) > $O/t9_tests.log 2>&1
echo "tests rc=$?"; tail -6 $O/t9_tests.log
timeout 300 python tools/pca_time.py 29600 0.0; SHOPT_PCA_NO_LIGHT=1 timeout 300 python tools/pca_time.py 29600 0.0
timeout 300 python tools/pca_time.py 14800 0.01
for c in c3 c5; do
  timeout 600 python bench.py --config $c --no-e2e > $O/t9_bench_$c.json 2> $O/t9_bench_$c.err
  python - <<PY
import json
d=json.load(open("$O/t9_bench_$c.json"))
st=d["roofline"]["stages"]
print("$c step %.2f ms value %.0f" % (d["ms_per_step"], d["value"]), {k:round(v["ms"],2) for k,v in st.items()}, d["mismatch_rate"])
PY
done

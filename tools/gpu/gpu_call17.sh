#!/bin/bash
# K = 21 with four pixels per thread in the ring kernels (SHOPT_POLY_PX4=1) against the default two
set -u
O=gpurun_out; mkdir -p $O
( SHOPT_POLY_PX4=1 timeout 900 python -m pytest tests/test_gpu_polyfit.py tests/test_gpu_baseline_shapes.py -x -q ) > $O/t17_tests_px4.log 2>&1
echo "tests rc=$?"; tail -3 $O/t17_tests_px4.log
for v in 0 1; do
  SHOPT_POLY_PX4=$v timeout 600 python bench.py --config c5 --no-cpu --no-e2e > $O/t17_bench_c5_px4_$v.json 2> $O/t17_bench_c5_px4_$v.err
  python - <<PY
import json
d=json.load(open("$O/t17_bench_c5_px4_$v.json"))
st=d["roofline"]["stages"]
print("c5 px4=$v step %.2f ms" % d["ms_per_step"], {k:(round(x["ms"],3), round(x["frac"],3)) for k,x in st.items() if "normal" in k or "mse" in k}, d["mismatch_rate"])
PY
done
echo done

#!/bin/bash
# normal_ring_kernel + mse_ring_kernel (SHOPT_POLY_RING=3): parity tests with the switch on, c3 / c5 / c4 stage times
set -u
O=gpurun_out; mkdir -p $O
( SHOPT_POLY_RING=3 timeout 900 python -m pytest tests/test_gpu_polyfit.py tests/test_gpu_stage.py tests/test_gpu_baseline_shapes.py tests/test_zz_no_alloc.py -x -q ) > $O/t15_tests_ring.log 2>&1
echo "tests rc=$?"; tail -3 $O/t15_tests_ring.log
for v in 3; do
  for c in c3 c5 c4; do
  SHOPT_POLY_RING=$v timeout 600 python bench.py --config $c --no-cpu --no-e2e > $O/t15_bench_${c}_ring$v.json 2> $O/t15_bench_${c}_ring$v.err
  python - <<PY
import json
d=json.load(open("$O/t15_bench_${c}_ring$v.json"))
st=d["roofline"]["stages"]
print("$c ring=$v step %.2f ms" % d["ms_per_step"], {k:(round(x["ms"],3), round(x["frac"],3)) for k,x in st.items() if "normal" in k or "mse" in k}, d["mismatch_rate"])
PY
  done
done
echo done

#!/bin/bash
# polyfit streaming kernels: parity, then A/B of the bulk-copy staging against plain loads on the bench workload
set -u
O=gpurun_out; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_polyfit.py tests/test_gpu_stage.py tests/test_zz_no_alloc.py -x -q ) > $O/t4_tests.log 2>&1
echo "tests rc=$?"; tail -4 $O/t4_tests.log
for bulk in 1 0; do
  SHOPT_POLY_BULK=$bulk timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/t4_bench_bulk$bulk.json 2> $O/t4_bench_bulk$bulk.err
  python - <<PY
import json
d=json.load(open("$O/t4_bench_bulk$bulk.json"))
st=d["roofline"]["stages"]
print("bulk=$bulk step %.2f ms" % d["ms_per_step"], {k:(round(v["ms"],3), round(v["frac"],3)) for k,v in st.items() if "normal" in k or "mse" in k})
PY
done

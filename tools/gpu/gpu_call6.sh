#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_polyfit.py tests/test_gpu_stage.py tests/test_gpu_baseline_shapes.py tests/test_gpu_modes.py -x -q ) > $O/t6_tests.log 2>&1
echo "tests rc=$?"; tail -4 $O/t6_tests.log
for c in c5 c3; do
  timeout 600 python bench.py --config $c > $O/t6_bench_$c.json 2> $O/t6_bench_$c.err
  python - <<PY
import json
d=json.load(open("$O/t6_bench_$c.json"))
st=d["roofline"]["stages"]
print("$c step %.2f ms value %.0f e2e %.0f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k:round(v["ms"],3) for k,v in st.items()}, d["mismatch_rate"])
PY
done

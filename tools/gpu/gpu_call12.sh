#!/bin/bash
# downdated polyfit rounds: whole GPU suite, c3 with the switch off/on, c5 and c4 lines
set -u
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/t12_gpu_tests.log 2>&1
echo "tests rc=$?"; tail -5 $O/t12_gpu_tests.log
SHOPT_POLY_DOWNDATE=0 timeout 600 python bench.py --config c3 --no-cpu > $O/t12_bench_c3_fresh.json 2> $O/t12_bench_c3_fresh.err
for c in c3 c4 c5; do
  timeout 600 python bench.py --config $c > $O/t12_bench_$c.json 2> $O/t12_bench_$c.err
done
for f in c3_fresh c3 c4 c5; do
  python - <<PY
import json
d=json.load(open("$O/t12_bench_$f.json"))
print("$f step %.2f ms value %.0f e2e %.0f launches %s" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"]), d["mismatch_rate"])
PY
done
SHOPT_TRACE=1 timeout 300 python bench.py --config c3 --steps 1 --warmup 1 --no-cpu --no-e2e > /dev/null 2> $O/t12_trace_c3.txt
grep -i "round\|rounds" $O/t12_trace_c3.txt | tail -8
echo done

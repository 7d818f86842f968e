#!/bin/bash
# last visit of round 2: whole GPU suite on the committed build, the c3 line once more
set -u
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/t21_gpu_tests.log 2>&1
echo "tests rc=$?"; tail -5 $O/t21_gpu_tests.log
timeout 600 python bench.py > $O/t21_bench_c3.json 2> $O/t21_bench_c3.err
python - <<PY
import json
d=json.load(open("$O/t21_bench_c3.json"))
st=d["roofline"]["stages"]
print("c3 step %.2f ms value %.0f e2e %.0f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k:(round(v["ms"],2), round(v["frac"],3), v.get("dram_bytes_per_launch")) for k,v in st.items()}, d["mismatch_rate"])
PY
echo done

#!/bin/bash
# fit_kernel pass A over pixel pairs (point symmetry of the profile): parity, then A/B
set -u
O=gpurun_out; mkdir -p $O
( timeout 1500 python -m pytest tests/test_gpu_analytic.py tests/test_gpu_stage.py tests/test_gpu_baseline_shapes.py tests/test_gpu_kolmogorov.py tests/test_gpu_pixelgrid.py -x -q ) > $O/t11_tests.log 2>&1
echo "tests rc=$?"; tail -6 $O/t11_tests.log
for sym in 1 0; do
  SHOPT_FIT_SYM=$sym timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e > $O/t11_bench_sym$sym.json 2> $O/t11_bench_sym$sym.err
  python - <<PY
import json
d=json.load(open("$O/t11_bench_sym$sym.json"))
st=d["roofline"]["stages"]
print("sym=$sym step %.2f ms" % d["ms_per_step"], {k:round(v["ms"],2) for k,v in st.items() if "fit" in k}, d["mismatch_rate"])
PY
done
SHOPT_FIT_SYM=1 timeout 300 python bench.py --config c2 --steps 10 --warmup 3 --no-e2e > $O/t11_c2.json 2>/dev/null; python -c "
import json;d=json.load(open('$O/t11_c2.json'));print('c2', d['ms_per_step'], d['mismatch_rate'])"

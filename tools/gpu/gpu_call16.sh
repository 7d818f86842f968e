#!/bin/bash
# pca_power_kernel with the cp.async column ring: parity tests (ring forced on for every stamp size), stage times
set -u
O=gpurun_out; mkdir -p $O
( SHOPT_PCA_RING=1 timeout 900 python -m pytest tests/test_gpu_pixelgrid.py tests/test_gpu_stage.py tests/test_gpu_baseline_shapes.py tests/test_gpu_modes.py -x -q ) > $O/t16_tests_ring.log 2>&1
echo "tests rc=$?"; tail -3 $O/t16_tests_ring.log
for v in 0 1; do
  for c in c3 c5; do
  SHOPT_PCA_RING=$v timeout 600 python bench.py --config $c --no-cpu --no-e2e > $O/t16_bench_${c}_ring$v.json 2> $O/t16_bench_${c}_ring$v.err
  python - <<PY
import json
d=json.load(open("$O/t16_bench_${c}_ring$v.json"))
st=d["roofline"]["stages"]
print("$c pca ring=$v step %.2f ms" % d["ms_per_step"], {k:(round(x["ms"],3), round(x["frac"],3)) for k,x in st.items() if "pca" in k}, d["mismatch_rate"])
PY
  done
done
for n in 3 5; do
SHOPT_PCA_LIGHT_CTAS=$n timeout 600 python bench.py --config c3 --no-cpu --no-e2e > $O/t16_bench_c3_ctas$n.json 2> $O/t16_bench_c3_ctas$n.err
python - <<PY
import json
d=json.load(open("$O/t16_bench_c3_ctas$n.json"))
st=d["roofline"]["stages"]
print("c3 ctas=$n step %.2f ms" % d["ms_per_step"], {k:(round(x["ms"],3), round(x["frac"],3)) for k,x in st.items() if "pca" in k})
PY
done
echo done

#!/bin/bash
# normal_ring_kernel (per-lane cp.async ring) against normal_kernel: parity tests with the switch on, c3 stage times, stalls
set -u
O=gpurun_out; mkdir -p $O
( SHOPT_POLY_RING=1 timeout 900 python -m pytest tests/test_gpu_polyfit.py tests/test_gpu_stage.py tests/test_gpu_baseline_shapes.py -x -q ) > $O/t14_tests_ring.log 2>&1
echo "tests rc=$?"; tail -3 $O/t14_tests_ring.log
for v in 0 1; do
  SHOPT_POLY_RING=$v timeout 600 python bench.py --config c3 --no-cpu --no-e2e > $O/t14_bench_c3_ring$v.json 2> $O/t14_bench_c3_ring$v.err
  SHOPT_POLY_RING=$v timeout 600 python bench.py --config c5 --no-cpu --no-e2e > $O/t14_bench_c5_ring$v.json 2> $O/t14_bench_c5_ring$v.err
  for c in c3 c5; do python - <<PY
import json
d=json.load(open("$O/t14_bench_${c}_ring$v.json"))
st=d["roofline"]["stages"]
print("$c ring=$v step %.2f ms" % d["ms_per_step"], {k:(round(x["ms"],3), round(x["frac"],3)) for k,x in st.items() if "normal" in k or "mse" in k})
PY
  done
done
SHOPT_POLY_RING=1 timeout 600 ncu --set full --clock-control none -k regex:'normal_ring_kernel' -c 1 -o /tmp/ring -f python bench.py --config c3 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t14_ncu.log 2>&1
ncu -i /tmp/ring.ncu-rep --page raw --csv > $O/t14_ring_raw.csv 2>> $O/t14_ncu.log
echo done

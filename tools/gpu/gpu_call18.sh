#!/bin/bash
# final build of round 2 (downdated rounds, ring kernels): whole GPU suite, bench lines of every config, launch list + full capture of c3
set -u
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/t18_gpu_tests.log 2>&1
echo "tests rc=$?"; tail -5 $O/t18_gpu_tests.log
for c in c3 c2 c4 c5; do
  extra=""; [ $c = c2 ] && extra="--steps 10 --warmup 3"
  timeout 600 python bench.py --config $c $extra > $O/t18_bench_$c.json 2> $O/t18_bench_$c.err
  python - <<PY
import json
d=json.load(open("$O/t18_bench_$c.json"))
st=d["roofline"]["stages"]
print("$c step %.2f ms value %.0f e2e %.0f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k:round(v["ms"],2) for k,v in st.items()}, d["mismatch_rate"])
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/t18_launches_c3.csv python bench.py --stars 20000 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t18_ncu_launches.log 2>&1
timeout 1200 ncu --set full --clock-control none -k regex:'fit_kernel|pca_power_kernel|pca_kernel|normal_ring_kernel|mse_ring_kernel|solve_kernel|sub_kernel' -c 14 -o /tmp/full_c3 -f python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t18_ncu_full.log 2>&1
ncu -i /tmp/full_c3.ncu-rep --page raw --csv > $O/t18_full_c3_raw.csv 2>> $O/t18_ncu_full.log
echo done

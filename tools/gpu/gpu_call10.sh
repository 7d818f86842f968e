#!/bin/bash
# after the two-stage PCA: whole GPU suite, bench lines of every config, launch list + full capture of c3
set -u
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/t10_gpu_tests.log 2>&1
echo "tests rc=$?"; tail -5 $O/t10_gpu_tests.log
for c in c3 c2 c4 c5; do
  extra=""; [ $c = c2 ] && extra="--steps 10 --warmup 3"
  timeout 600 python bench.py --config $c $extra > $O/t10_bench_$c.json 2> $O/t10_bench_$c.err
  python - <<PY
import json
d=json.load(open("$O/t10_bench_$c.json"))
st=d["roofline"]["stages"]
print("$c step %.2f ms value %.0f e2e %.0f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k:round(v["ms"],2) for k,v in st.items()}, d["mismatch_rate"])
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/t10_launches_c3.csv python bench.py --stars 20000 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t10_ncu_launches.log 2>&1
timeout 1200 ncu --set full --clock-control none -k regex:'fit_kernel|pca_power_kernel|pca_kernel|normal_kernel|mse_kernel|solve_kernel' -c 12 -o /tmp/full_c3 -f python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t10_ncu_full.log 2>&1
ncu -i /tmp/full_c3.ncu-rep --page raw --csv > $O/t10_full_c3_raw.csv 2>> $O/t10_ncu_full.log
echo done

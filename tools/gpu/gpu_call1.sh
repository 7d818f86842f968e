#!/bin/bash
# One GPU box visit: parity tests, bench lines of every config, ncu launch list and one --set full capture.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $O/smi.txt 2>&1
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/gpu_tests.log 2>&1
echo "tests rc=$?" >> $O/gpu_tests.log
timeout 600 python bench.py > $O/bench_c3_n1.json 2> $O/bench_c3_n1.err
echo "bench c3 rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_ref.json 2> $O/bench_c3_ref.err
timeout 400 python bench.py --config c2 --steps 10 --warmup 3 > $O/bench_c2_n1.json 2> $O/bench_c2_n1.err
echo "bench c2 rc=$?"
timeout 600 python bench.py --config c4 > $O/bench_c4_n1.json 2> $O/bench_c4_n1.err
echo "bench c4 rc=$?"
timeout 600 python bench.py --config c5 > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err
echo "bench c5 rc=$?"
# launch list (cold-cache, serialised: compare shares)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv \
  python bench.py --stars 20000 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
# one full capture of the step's kernels at the bench size (first complete step: fit, pca, fit, 3 x (normal, solve.., mse))
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'fit_kernel|pca_kernel|normal_kernel|mse_kernel|solve_kernel' -c 9 \
  -o /tmp/full_c3 -f python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > $O/ncu_full.log 2>&1
echo "full capture rc=$?"
# the report itself is too large to travel (64 MiB cap on gpurun_out): export the pages here
ncu -i /tmp/full_c3.ncu-rep --page raw --csv > $O/full_c3_raw.csv 2>> $O/ncu_full.log
ncu -i /tmp/full_c3.ncu-rep --page source --csv -k regex:fit_kernel -c 1 > $O/full_c3_source_fit.csv 2>> $O/ncu_full.log
ncu -i /tmp/full_c3.ncu-rep --page source --csv -k regex:pca_kernel -c 1 > $O/full_c3_source_pca.csv 2>> $O/ncu_full.log
ncu -i /tmp/full_c3.ncu-rep --page source --csv -k regex:normal_kernel -c 1 > $O/full_c3_source_normal.csv 2>> $O/ncu_full.log
ncu -i /tmp/full_c3.ncu-rep --page source --csv -k regex:mse_kernel -c 1 > $O/full_c3_source_mse.csv 2>> $O/ncu_full.log
ls -la /tmp/full_c3.ncu-rep; du -sh $O
ls -la $O

#!/bin/bash
# ncu evidence for the K = 21 decision (c5: 301 x 301, degree 5) and for the masked catalogue (c4)
set -u
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none -k regex:'normal_kernel|mse_kernel|solve_kernel|factor_kernel' -c 8 \
  -o /tmp/full_c5 -f python bench.py --config c5 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t5_ncu_c5.log 2>&1
echo "c5 capture rc=$?"
ncu -i /tmp/full_c5.ncu-rep --page raw --csv > $O/full_c5_raw.csv 2>> $O/t5_ncu_c5.log
timeout 900 ncu --set full --clock-control none -k regex:'pca_kernel' -c 1 \
  -o /tmp/full_c4 -f python bench.py --config c4 --stars 14800 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t5_ncu_c4.log 2>&1
echo "c4 capture rc=$?"
ncu -i /tmp/full_c4.ncu-rep --page raw --csv > $O/full_c4_pca_raw.csv 2>> $O/t5_ncu_c4.log
timeout 600 python bench.py --config c4 > $O/t5_bench_c4_n1.json 2> $O/t5_bench_c4_n1.err; echo "c4 bench rc=$?"
du -sh $O

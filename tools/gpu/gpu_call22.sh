#!/bin/bash
# ncu --set full of the polyfit ring kernels at K = 21 (c5) for the K = 21 paragraph of DESIGN.md 4.3
set -u
O=gpurun_out; mkdir -p $O
timeout 500 ncu --set full --clock-control none -k regex:'normal_ring_kernel|mse_ring_kernel' -c 2 -o /tmp/c5ring -f python bench.py --config c5 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/t22_ncu.log 2>&1
ncu -i /tmp/c5ring.ncu-rep --page raw --csv > $O/t22_c5_ring_raw.csv 2>> $O/t22_ncu.log
wc -c $O/t22_c5_ring_raw.csv
echo done

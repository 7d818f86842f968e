#!/bin/bash
# 2-GPU box: the sharded stage against the single-GPU stage (incl. median_constraint and the NaN repair), then the
# whole GPU suite and a 2-GPU bench line
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/t3_smi.txt 2>&1
( timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q -s ) > $O/t3_multirank.log 2>&1
echo "multirank rc=$?"; tail -15 $O/t3_multirank.log
( timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/t3_gpu_tests.log 2>&1
echo "gpu tests rc=$?"; tail -4 $O/t3_gpu_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 2 --steps 3 --warmup 3 > $O/t3_bench_c3_n2.json 2> $O/t3_bench_c3_n2.err
echo "bench n2 rc=$?"; cut -c1-400 $O/t3_bench_c3_n2.json

#!/bin/bash
# 8-GPU box: the BASELINE configurations at their stated total sizes (c3 800k, c4 1M masked, c5 50k x 301^2), weak scaling
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/t7_smi.txt 2>&1
N=$(nvidia-smi -L | wc -l)
run() { # config name
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N --config $1 --steps 3 --warmup 3 > $O/t7_bench_$1_n$N.json 2> $O/t7_bench_$1_n$N.err
  echo "$1 rc=$?"; cut -c1-260 $O/t7_bench_$1_n$N.json
}
run c3 29701
run c4 29702
run c5 29703
( timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -s ) > $O/t7_multirank.log 2>&1; echo "multirank rc=$?"; tail -3 $O/t7_multirank.log

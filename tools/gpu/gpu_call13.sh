#!/bin/bash
# N-GPU box with the final build: the sharded-vs-single-GPU worker (covers the sharded downdate) and the c3 line; with
# ALL=1 also c4 and c5 at N GPUs
set -u
O=gpurun_out; mkdir -p $O
N=$(nvidia-smi -L | wc -l)
( timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -s ) > $O/t13_multirank_n$N.log 2>&1; echo "multirank rc=$?"; grep "MULTIRANK_OK\|passed\|failed" $O/t13_multirank_n$N.log | tail -3
run() { # config port
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N --config $1 --steps 3 --warmup 3 > $O/t13_bench_$1_n$N.json 2> $O/t13_bench_$1_n$N.err
  echo "$1 rc=$?"; cut -c1-200 $O/t13_bench_$1_n$N.json
}
run c3 29711
if [ "${ALL:-0}" = 1 ]; then run c4 29712; run c5 29713; fi
echo done

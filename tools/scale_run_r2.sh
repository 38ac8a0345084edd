#!/bin/bash
# Round-2 scaling session on one 8xB200 box: the 1M-entry cache (strong scaling, the headline
# config) and the 10M-entry cache of BASELINE.json configs[2] at N = 1, 2, 4, 8.
mkdir -p gpurun_out
run() { # N entries steps warmup tag extra...
  N=$1; E=$2; K=$3; W=$4; TAG=$5; shift 5
  if [ $N = 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps $K --warmup $W --entries $E --no-extra --no-cpu-baseline "$@" \
      >> gpurun_out/r2_scale_$TAG.jsonl 2>> gpurun_out/r2_scale.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600 + N)) bench.py --gpus $N --steps $K --warmup $W --entries $E "$@" \
      >> gpurun_out/r2_scale_$TAG.jsonl 2>> gpurun_out/r2_scale.err
  fi
}
for N in 1 2 4 8; do run $N 1000000 20 5 1m; done
for N in 1 2 4 8; do run $N 1000000 200 10 1m_200; done
for N in 8 4 2 1; do run $N 10000000 10 3 10m --oracle-queries 8; done
run 8 1000000 20 5 1m_noslice --no-slices
run 8 1000000 20 5 1m_serial --no-overlap
run 8 1000000 20 5 1m_rsv4 --reserve-sms 4

#!/bin/bash
# Multi-GPU scaling session (run under `gpurun --gpus 8`): strong scaling of the 1M-entry headline
# workload at N = 2, 4, 8 (pipelined vs serial exchange) and BASELINE config 3 (10M entries) at
# N = 1 and 8.  One JSON line per run is appended to gpurun_out/scale.jsonl.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/scale.jsonl
: > $OUT
run() {  # n, extra args...
  n=$1; shift
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --no-cpu-baseline "$@" 2>>gpurun_out/scale.err | tail -1 >> $OUT
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" 2>>gpurun_out/scale.err | tail -1 >> $OUT
  fi
  echo "run n=$n $* -> rc=$?"
}
S="--steps 200 --warmup 10"
run 1 $S
run 2 $S --no-overlap
run 4 $S --no-overlap
run 4 $S --reserve-sms 4
run 8 $S --no-overlap
run 8 $S --reserve-sms 4
run 8 $S --reserve-sms 8
run 8 --steps 100 --warmup 5 --entries 10000000 --reserve-sms 4
run 1 --steps 30 --warmup 3 --entries 10000000
python - <<'PY'
import json
for l in open("gpurun_out/scale.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:100]); continue
    print(d["n_gpus"], d["config"]["cache_entries"], round(d["value"]), round(d["ms_per_step"], 4),
          round(d["roofline"]["launch_ms"], 4), d["clocks"]["sm_mhz"], d["config"]["parallelism"][:40])
PY

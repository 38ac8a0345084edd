#!/bin/bash
# Final round-1 scaling table with the default settings (plus N=2 peer-exchange variants).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/scale3.jsonl
: > $OUT
run() {  # label, n, extra args...
  label=$1; n=$2; shift 2
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --no-cpu-baseline "$@" 2>>gpurun_out/scale3.err | tail -1 | sed "s/^{/{\"label\": \"$label\", /" >> $OUT
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" 2>>gpurun_out/scale3.err | tail -1 | sed "s/^{/{\"label\": \"$label\", /" >> $OUT
  fi
  echo "run $label n=$n $* -> rc=$?"
}
S="--steps 300 --warmup 10"
run default 1 $S
run default 2 $S
run p2p_serial 2 $S --exchange p2p
run default 4 $S
run default 8 $S
run cfg3 8 --steps 100 --warmup 5 --entries 10000000
python - <<'PY'
import json
for l in open("gpurun_out/scale3.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:120]); continue
    print(d["label"], d["n_gpus"], d["config"]["cache_entries"], round(d["value"]), round(d["ms_per_step"], 4),
          round(d["roofline"]["launch_ms"], 4), round(d["e2e"]["value"]), d["clocks"].get("sm_mhz"), d["config"]["parallelism"][:90])
PY

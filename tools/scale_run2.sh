#!/bin/bash
# Second multi-GPU session: final kernels (CTA-pair scan), pipelined e2e leg, NCCL channel limits.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/scale2.jsonl
: > $OUT
run() {  # label, n, extra args...
  label=$1; n=$2; shift 2
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --no-cpu-baseline "$@" 2>>gpurun_out/scale2.err | tail -1 | sed "s/^{/{\"label\": \"$label\", /" >> $OUT
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" 2>>gpurun_out/scale2.err | tail -1 | sed "s/^{/{\"label\": \"$label\", /" >> $OUT
  fi
  echo "run $label n=$n $* -> rc=$?"
}
S="--steps 300 --warmup 10"
run base 1 $S
run base 2 $S
run base 4 $S
run base 8 $S
run serial 8 $S --no-overlap
NCCL_MAX_NCHANNELS=2 run nch2 8 $S
NCCL_MAX_NCHANNELS=4 run nch4 8 $S
NCCL_MAX_NCHANNELS=2 run nch2_r2 8 $S --reserve-sms 2
run cfg3 8 --steps 100 --warmup 5 --entries 10000000
python - <<'PY'
import json
for l in open("gpurun_out/scale2.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:120]); continue
    print(d["label"], d["n_gpus"], d["config"]["cache_entries"], round(d["value"]), round(d["ms_per_step"], 4),
          round(d["roofline"]["launch_ms"], 4), round(d["e2e"]["value"]), d["clocks"].get("sm_mhz"), d["clocks"].get("reasons"))
PY

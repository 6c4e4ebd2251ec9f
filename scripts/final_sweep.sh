#!/bin/bash
# Final measurement at N GPUs (run as: gpurun --gpus N -- 'bash scripts/final_sweep.sh N [ref]').
# Writes gpurun_out/final_n${N}_*.json. Weak scaling (per-GPU batch 64, BASELINE config 2) and
# strong scaling (global batch 512, config 3); optionally the reference arm.
N=${1:-1}
REF=${2:-}
mkdir -p gpurun_out
PORT=$((29800 + N))
run() {  # run <out> <args...>
  out=$1; shift
  if [ "$N" -gt 1 ]; then
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $PORT bench.py --gpus $N "$@" 2>gpurun_out/${out}.err | tail -n 1 > gpurun_out/${out}.json
  else
    timeout 300 python bench.py --gpus 1 "$@" 2>gpurun_out/${out}.err | tail -n 1 > gpurun_out/${out}.json
  fi
  PORT=$((PORT + 10))
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${out}.json").read())
    e = d.get("e2e") or {}
    print("${out}", "value", round(d.get("value") or 0), "ms/step", round(d.get("ms_per_step") or 0, 4),
          "e2e", round(e.get("value") or 0), d.get("clocks"))
except Exception as ex:
    print("${out}", "FAILED", ex, open("gpurun_out/${out}.err").read()[-800:])
PY
}
run final_n${N}_weak --steps 1876 --warmup 50
run final_n${N}_strong --steps 938 --warmup 30 --batch $((512 / N))
if [ -n "$REF" ]; then
  run final_n${N}_ref_weak --impl reference --steps 300 --warmup 10
  run final_n${N}_ref_strong --impl reference --steps 200 --warmup 10 --batch $((512 / N)) --ref-cudnn on
fi

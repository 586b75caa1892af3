#!/bin/bash
# Sweep of the chunked multi-GPU sort's knobs on N GPUs of one box (run under gpurun):
#   scripts/sweep_multi.sh N TAG "ENV1=a ENV2=b" "ENV1=c" ...
# Every configuration runs `bench.py --gpus N --no-e2e --no-cpu` with XCHG_TIMELINE=1 and
# leaves its JSON line and per-rank GPU timelines in gpurun_out/TAG_sweep.log.
N=$1; TAG=$2; shift 2
OUT=gpurun_out/${TAG}_sweep.log
: > $OUT
PORT=29510
for CFG in "$@"; do
  PORT=$((PORT+1))
  echo "=== $CFG" >> $OUT
  env $CFG XCHG_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
      --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-cpu $SWEEP_EXTRA \
      > gpurun_out/${TAG}_tmp.json 2> gpurun_out/${TAG}_tmp.err
  echo "rc=$?" >> $OUT
  grep -E "gpu timeline rank 0|Error|error" gpurun_out/${TAG}_tmp.err | tail -4 >> $OUT
  python - >> $OUT <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_tmp.json").read().strip().splitlines()[-1])
    print("ms_per_step", round(d["ms_per_step"], 3), "value", round(d["value"] / 1e9, 3), "G rows/s", "verified", d.get("verified"))
    print("kernels", [(k["kernel"], round(k["ms_per_step"], 2), k["launches_per_step"]) for k in d.get("kernels", [])][:6])
except Exception as e:
    print("no json line:", e)
PY
done
cat $OUT

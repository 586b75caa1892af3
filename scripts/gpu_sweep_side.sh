#!/bin/bash
for cap in 1 2 4 0; do
  XCHG_SIDE_CTAS_PER_SM=$cap timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/side_$cap.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/side_$cap.log") if x.startswith("{")]
if not l: print("cap=$cap FAILED", open("gpurun_out/side_$cap.log").read()[-1500:])
else:
    d=json.loads(l[-1]); print("cap=$cap ms/step", round(d["ms_per_step"],2), [(k["kernel"], round(k["ms_per_step"],2)) for k in d["kernels"][:3]])
PY
done
XCHG_OVERLAP=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/side_off.log 2>&1
grep -o '"ms_per_step": [0-9.]*' gpurun_out/side_off.log | head -1
python scripts/bench_ops.py --steps 3 --warmup 2 > gpurun_out/ops.log 2>&1; grep "^{" gpurun_out/ops.log | cut -c1-900; tail -3 gpurun_out/ops.log | grep -v "^{" | cut -c1-600

#!/bin/bash
# GPU box helper (2 GPUs): all GPU tests, N=1 bench with e2e, N=2 bench; prints summaries.
tag=${1:-check2}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/${tag}_bench1.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/${tag}_bench2.log 2>&1
echo "bench2 rc=$?"
python - <<PY
import json
for f in ("gpurun_out/${tag}_bench1.log", "gpurun_out/${tag}_bench2.log"):
    l=[x for x in open(f) if x.startswith("{")]
    if not l: print(f, open(f).read()[-3000:]); continue
    d=json.loads(l[-1]); print(f, "ms/step", round(d["ms_per_step"],2), "value", round(d["value"]/1e9,3), "Grows/s", [(k["kernel"], round(k["ms_per_step"],2), k["launches_per_step"]) for k in d["kernels"][:4]]); print("   e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"]/1e6, "Mrows/s")
PY

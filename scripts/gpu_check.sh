#!/bin/bash
# GPU box helper: sort-path tests, then the N=1 bench (with e2e and phase trace); prints a summary.
tag=${1:-check}
python -m pytest tests/test_sort_parity.py tests/test_device_exchange.py tests/test_exchange_kernels.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-cpu --trace > gpurun_out/${tag}_bench.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/${tag}_bench.log") if x.startswith("{")]
if not l: print(open("gpurun_out/${tag}_bench.log").read()[-2000:])
else:
    d=json.loads(l[-1]); print("ms/step", round(d["ms_per_step"],2), [(k["kernel"], round(k["ms_per_step"],2), k["launches_per_step"]) for k in d["kernels"][:4]]); print(d["e2e"])
PY
grep "trace rank" gpurun_out/${tag}_bench.log

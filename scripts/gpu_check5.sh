#!/bin/bash
timeout 600 python -m pytest tests/test_radix_sort_kernel.py tests/test_sort_parity.py -m gpu -x -q 2>&1 | tail -5
for algo in lsd msd; do
XCHG_SORT_ALGO=$algo python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/b_$algo.log 2>&1; grep "^{" gpurun_out/b_$algo.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$algo ms/step', d['ms_per_step'], [(k['kernel'],k['launches_per_step'],round(k['ms_per_step'],3)) for k in d['kernels']])"
done

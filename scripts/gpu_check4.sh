#!/bin/bash
# new sort kernel tests first, then the whole GPU suite, then A/B bench lines
timeout 600 python -m pytest tests/test_radix_sort_kernel.py -x -q 2>&1 | tail -15
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
XCHG_SORT_ALGO=lsd python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/b_lsd.log 2>&1; grep "^{" gpurun_out/b_lsd.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('LSD ms/step', d['ms_per_step'], [(k['kernel'],k['launches_per_step'],round(k['ms_per_step'],3)) for k in d['kernels']])"
python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/b_msd.log 2>&1; grep "^{" gpurun_out/b_msd.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('MSD ms/step', d['ms_per_step'], [(k['kernel'],k['launches_per_step'],round(k['ms_per_step'],3)) for k in d['kernels']])"
tail -5 gpurun_out/b_msd.log | cut -c1-600
XCHG_LS_WIDE2=1 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/b_msd_w2.log 2>&1; grep "^{" gpurun_out/b_msd_w2.log | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('MSD wide2 ms/step', d['ms_per_step'], [(k['kernel'],k['launches_per_step'],round(k['ms_per_step'],3)) for k in d['kernels']])"

#!/bin/bash
# round-1 final profiles: launch list + one full capture of the step's main kernels
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_r1f.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1f.csv $CMD > gpurun_out/ncu_list_r1f.log 2>&1
echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_gather_rows|k_onesweep|k_local_sort|k_normalize_hist|k_seg_hist" -s 18 -c 6 -o gpurun_out/prof_r1f $CMD > gpurun_out/ncu_full_r1f.log 2>&1
echo "full rc=$?"
grep "^{" gpurun_out/plain_r1f.log | cut -c1-200

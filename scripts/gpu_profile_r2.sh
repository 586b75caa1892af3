#!/bin/bash
# round-2 profiles of the one-GPU step: plain run, launch list, one full capture of the
# step's main kernels (run under gpurun on one GPU; the .ncu-rep is read back here with
# scripts/ncu_summary.py and summarised into profiles/)
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-ops"
$CMD > gpurun_out/r02_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1
echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_gather_rows|k_onesweep|k_local_sort|k_key_scan|k_seg_hist|k_denormalize" -s 21 -c 7 -o gpurun_out/r02_prof $CMD > gpurun_out/r02_ncu_full.log 2>&1
echo "full rc=$?"
grep "^{" gpurun_out/r02_plain.log | cut -c1-200

#!/bin/bash
# compute-sanitizer over the kernel-level GPU tests (small sizes): memcheck (out-of-bounds /
# misaligned global and shared accesses) and racecheck (shared-memory hazards: the sort's
# warp-private counters ordered by __syncwarp, the aliased shared regions of k_onesweep and
# k_local_sort).  Summaries land in gpurun_out/ (copy what is to be kept into profiles/).
TESTS="tests/test_radix_sort_kernel.py tests/test_exchange_control.py tests/test_merge.py tests/test_exchange_kernels.py tests/test_shuffle.py"
TAG=${1:-r02}
for TOOL in memcheck racecheck; do
  OUT=gpurun_out/${TAG}_sanitizer_${TOOL}.log
  timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 9 --print-limit 20 \
      python -m pytest $TESTS -m gpu -x -q -p no:cacheprovider > $OUT 2>&1
  echo "$TOOL rc=$?" | tee -a $OUT
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" $OUT | tail -5
done

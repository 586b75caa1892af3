#!/bin/bash
# A/B of the row gather for 16-byte-aligned rows (random_shuffle of 1 KiB tensor rows):
# cp.async.cg staging + per-thread stores (default) vs cp.async.bulk through the TMA unit.
for MODE in 0 1; do
  echo "=== XCHG_GATHER_BULK=$MODE"
  XCHG_GATHER_BULK=$MODE python scripts/bench_ops.py --only shuffle --steps 10 --warmup 3
done

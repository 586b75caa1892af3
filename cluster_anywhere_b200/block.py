"""Block metadata types of the operator surface (python/ray/data/block.py:141-252)."""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import List, Optional

import pyarrow as pa


@dataclass
class BlockExecStats:
    """Per-task execution stats (python/ray/data/block.py:141-207).  ``gpu_time_s``
    is this engine's addition: device time measured with CUDA events."""

    start_time_s: Optional[float] = None
    end_time_s: Optional[float] = None
    wall_time_s: Optional[float] = None
    cpu_time_s: Optional[float] = None
    gpu_time_s: Optional[float] = None
    node_id: str = "b200"

    @staticmethod
    def builder() -> "_StatsBuilder":
        return _StatsBuilder()


class _StatsBuilder:
    def __init__(self):
        self._t0 = time.perf_counter()
        self._c0 = time.process_time()

    def build(self) -> BlockExecStats:
        t1, c1 = time.perf_counter(), time.process_time()
        return BlockExecStats(self._t0, t1, t1 - self._t0, c1 - self._c0)


@dataclass
class BlockMetadata:
    """python/ray/data/block.py:234-252."""

    num_rows: Optional[int]
    size_bytes: Optional[int]
    schema: Optional[pa.Schema]
    input_files: Optional[List[str]] = None
    exec_stats: Optional[BlockExecStats] = None


def empty_table() -> pa.Table:
    """The schema-less empty block the reference hands out for empty inputs
    (python/ray/data/_internal/arrow_block.py:316, :445-446)."""
    return pa.table({})

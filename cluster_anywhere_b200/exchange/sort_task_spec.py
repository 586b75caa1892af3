"""``SortTaskSpec`` on the GPU (python/ray/data/_internal/planner/exchange/sort_task_spec.py:92-230)."""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np
import pyarrow as pa

from ..block import BlockExecStats, BlockMetadata
from ..block_accessor import B200ArrowBlockAccessor
from ..sort_key import SortKey
from .interfaces import ExchangeTaskSpec


class _Last:
    """Sorts after every value: stands in for None / NaN samples
    (the reference's NULL_SENTINEL, python/ray/data/_internal/util.py:67-89)."""

    def __lt__(self, other):
        return False

    def __gt__(self, other):
        return not isinstance(other, _Last)

    def __eq__(self, other):
        return isinstance(other, _Last)

    def __le__(self, other):
        return isinstance(other, _Last)

    def __ge__(self, other):
        return True

    def __hash__(self):
        return 0


_LAST = _Last()


def _is_na(x) -> bool:
    if x is None:
        return True
    try:
        return bool(isinstance(x, (float, np.floating)) and np.isnan(x))
    except TypeError:
        return False


class SortTaskSpec(ExchangeTaskSpec):
    """sample -> sort+partition every block -> merge each key range."""

    SORT_SAMPLE_SUB_PROGRESS_BAR_NAME = "Sort Sample"

    def __init__(self, boundaries: list, sort_key: SortKey, batch_format: str = "default"):
        super().__init__(map_args=[boundaries, sort_key], reduce_args=[sort_key, batch_format])

    @staticmethod
    def map(idx: int, block, output_num_blocks: int, boundaries: list, sort_key: SortKey) -> list:
        """sort_task_spec.py:127-138: R range partitions of the sorted block, metadata last."""
        stats = BlockExecStats.builder()
        acc = B200ArrowBlockAccessor.for_block(block)
        out = acc.sort_and_partition(boundaries, sort_key)
        meta = acc.get_metadata(exec_stats=stats.build())
        return out + [meta]

    @staticmethod
    def reduce(sort_key: SortKey, batch_format: str, *mapper_outputs,
               partial_reduce: bool = False) -> Tuple[pa.Table, BlockMetadata]:
        """sort_task_spec.py:140-152: merge one partition from every mapper."""
        from ..block_accessor import B200PandasBlockAccessor, want_pandas_output

        cls = B200PandasBlockAccessor if want_pandas_output(mapper_outputs, batch_format) else B200ArrowBlockAccessor
        blocks = [B200ArrowBlockAccessor.for_block(b).to_arrow() for b in mapper_outputs]
        return cls.merge_sorted_blocks(blocks, sort_key)

    @staticmethod
    def boundaries_from_samples(samples: List[tuple], sort_key: SortKey, num_reducers: int) -> list:
        """Driver half of ``sample_boundaries`` (sort_task_spec.py:185-226)."""
        if len(samples) == 0:
            return [None] * (num_reducers - 1)
        s = list(samples)
        for i, desc in list(enumerate(sort_key.get_descending()))[::-1]:
            s.sort(key=lambda t, i=i: _LAST if _is_na(t[i]) else t[i], reverse=desc)
        q = [int(x * (len(s) - 1)) for x in np.linspace(0, 1, num_reducers + 1)]
        return [s[i] for i in q[1:-1]]

    @staticmethod
    def sample_boundaries(blocks: List[pa.Table], sort_key: SortKey, num_reducers: int,
                          sample_bar=None) -> list:
        """sort_task_spec.py:154-226: ``int(R*10/M)`` key samples per block, then the
        R-1 quantile tuples.  Blocks are host Arrow tables here (the reference
        takes ObjectRefs and runs one task per block)."""
        columns = sort_key.get_columns()
        n_samples = int(num_reducers * 10 / len(blocks))
        tables = [B200ArrowBlockAccessor.for_block(b).sample(n_samples, sort_key) for b in blocks]
        tables = [t for t in tables if t.num_rows > 0]
        if not tables:
            return [None] * (num_reducers - 1)
        merged = pa.concat_tables(tables, promote_options="permissive")
        cols = [merged[c].to_pylist() for c in columns]
        return SortTaskSpec.boundaries_from_samples(list(zip(*cols)), sort_key, num_reducers)

"""``SortAggregateTaskSpec`` on the GPU
(python/ray/data/_internal/planner/exchange/aggregate_task_spec.py:10-109)."""
from __future__ import annotations

from typing import List, Tuple

import pyarrow as pa

from .. import aggregate as agg
from .. import device as dev
from ..block import BlockExecStats, BlockMetadata, empty_table
from ..block_accessor import B200ArrowBlockAccessor
from ..sort_key import SortKey
from .interfaces import ExchangeTaskSpec


class SortAggregateTaskSpec(ExchangeTaskSpec):
    """map: partial aggregate of one block, cut into the R key ranges; reduce: merge the
    partials of one key range and finalize."""

    def __init__(self, boundaries: list, key: SortKey, aggs: list, batch_format: str = "default"):
        super().__init__(map_args=[boundaries, key, aggs], reduce_args=[key, aggs, batch_format])

    @staticmethod
    def map(idx: int, block, output_num_blocks: int, boundaries: list, sort_key: SortKey,
            aggs: list) -> list:
        """aggregate_task_spec.py:38-59.  The reference sorts the block, slices it at the
        boundaries and combines every slice; the groups of different slices are
        disjoint key ranges, so ONE group aggregate of the whole block (K8, output
        sorted by key) cut at the same boundaries (K4) yields the same R partial
        blocks."""
        stats = BlockExecStats.builder()
        acc = B200ArrowBlockAccessor.for_block(block)
        table = agg.prune_columns(acc.to_arrow(), sort_key.get_columns(), aggs)
        acc = B200ArrowBlockAccessor(table)
        key = B200ArrowBlockAccessor._group_key(sort_key)
        if key is None:
            parts = [acc.combine(sort_key, aggs)]
        elif acc.num_rows() == 0:
            parts = [empty_table() for _ in range(len(boundaries) + 1)]
        else:
            rt = acc.rt
            out = agg.aggregate_table(rt, dev.stage_table(rt, table), key, aggs, finalize=False)
            cuts = agg.group_cut_points(rt, out, key, boundaries)
            host = dev.export_table(rt, out)
            parts, last = [], 0
            for c in cuts:
                parts.append(host[last:c])
                last = c
            parts.append(host[last:])
        return parts + [acc.get_metadata(exec_stats=stats.build())]

    @staticmethod
    def reduce(key: SortKey, aggs: list, batch_format: str, *mapper_outputs,
               partial_reduce: bool = False) -> Tuple[pa.Table, BlockMetadata]:
        """aggregate_task_spec.py:61-74."""
        from ..block_accessor import B200PandasBlockAccessor

        blocks = [B200ArrowBlockAccessor.for_block(b).to_arrow() for b in mapper_outputs]
        # partial blocks are the engine's own format; only a FINAL block follows batch_format
        cls = B200PandasBlockAccessor if (batch_format == "pandas" and not partial_reduce) else B200ArrowBlockAccessor
        return cls.aggregate_combined_blocks(blocks, key, aggs, finalize=not partial_reduce)

    @staticmethod
    def _prune_unused_columns(block, sort_key: SortKey, aggs):
        return agg.prune_columns(B200ArrowBlockAccessor.for_block(block).to_arrow(),
                                 sort_key.get_columns(), aggs)

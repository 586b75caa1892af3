"""User-level group-by surface over in-memory blocks: ``GroupedData`` and the
multi-column aggregate builder.

Mirrors ``GroupedData.aggregate`` / ``_aggregate_on`` and the ``count / sum / min / max /
mean / std`` shortcuts (python/ray/data/grouped_data.py:50-93, :270-560) and
``Dataset._build_multicolumn_aggs`` (python/ray/data/dataset.py:5362-5388).  The reference
builds an ``Aggregate`` logical operator and lets its planner run it lazily; here the
blocks are in memory, so ``aggregate`` plans the same ``Aggregate`` operator with this
package's planner (``planner.plan_all_to_all_op`` -> ``generate_aggregate_fn`` ->
``SortAggregateTaskSpec`` on the GPU, or the whole operator on the device under
``ShuffleStrategy.B200_DEVICE_EXCHANGE``) and returns the output blocks.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Union

import pyarrow as pa

from . import planner as P
from .aggregate import AggregateFn, Count, Max, Mean, Min, Std, Sum


def build_multicolumn_aggs(schema: Optional[pa.Schema], agg_cls: type, on: Optional[Union[str, List[str]]],
                           *args, skip_cols: Optional[Union[str, List[str]]] = None, **kwargs) -> list:
    """One aggregation per column (dataset.py:5362-5388): ``on=None`` expands to every
    column of the schema that is not a group key; a single name stays a single aggregation."""
    if on is None and schema is not None and len(schema.names) > 0:
        skip = [] if not skip_cols else ([skip_cols] if isinstance(skip_cols, str) else list(skip_cols))
        on = [c for c in schema.names if c not in skip]
    if not isinstance(on, list):
        on = [on]
    if len(on) == 0:
        raise ValueError("At least 1 column to aggregate on has to be provided")
    return [agg_cls(on_, *args, **kwargs) for on_ in on]


class GroupedData:
    """Blocks grouped by ``key`` (a column name, a list of names, or None = one global
    group); the grouping happens when an aggregation is applied."""

    def __init__(self, blocks: Sequence[pa.Table], key: Optional[Union[str, List[str]]], *,
                 num_partitions: Optional[int] = None, data_context: Optional[P.DataContext] = None):
        self._blocks = list(blocks)
        self._key = key
        self._num_partitions = num_partitions
        self._ctx = data_context

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}(blocks={len(self._blocks)}, key={self._key!r})"

    def _schema(self) -> Optional[pa.Schema]:
        for b in self._blocks:
            if b.num_columns > 0:
                return b.schema
        return None

    def aggregate(self, *aggs: AggregateFn) -> List[pa.Table]:
        """grouped_data.py:50-74: key column(s) first, then one column per aggregation;
        without a key the result is a single row."""
        op = P.Aggregate(P.InputData(), self._key, list(aggs), num_partitions=self._num_partitions)
        bundles, _ = P.execute_all_to_all(op, [P.RefBundle.from_block(b) for b in self._blocks], self._ctx)
        return [blk for rb in bundles for blk in rb.block_refs]

    def _aggregate_on(self, agg_cls: type, on: Optional[Union[str, List[str]]], *args, **kwargs) -> List[pa.Table]:
        """grouped_data.py:76-93."""
        aggs = build_multicolumn_aggs(self._schema(), agg_cls, on, *args, skip_cols=self._key, **kwargs)
        return self.aggregate(*aggs)

    # the shortcuts of grouped_data.py:270-560
    def count(self) -> List[pa.Table]:
        return self.aggregate(Count())

    def sum(self, on=None, ignore_nulls: bool = True) -> List[pa.Table]:
        return self._aggregate_on(Sum, on, ignore_nulls=ignore_nulls)

    def min(self, on=None, ignore_nulls: bool = True) -> List[pa.Table]:
        return self._aggregate_on(Min, on, ignore_nulls=ignore_nulls)

    def max(self, on=None, ignore_nulls: bool = True) -> List[pa.Table]:
        return self._aggregate_on(Max, on, ignore_nulls=ignore_nulls)

    def mean(self, on=None, ignore_nulls: bool = True) -> List[pa.Table]:
        return self._aggregate_on(Mean, on, ignore_nulls=ignore_nulls)

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True) -> List[pa.Table]:
        return self._aggregate_on(Std, on, ignore_nulls=ignore_nulls, ddof=ddof)


def groupby(blocks: Sequence[pa.Table], key: Optional[Union[str, List[str]]],
            num_partitions: Optional[int] = None, data_context: Optional[P.DataContext] = None) -> GroupedData:
    """``Dataset.groupby`` (python/ray/data/dataset.py:2119-2180) for blocks already in memory."""
    if key is not None:
        from .sort_key import SortKey

        schema = next((b.schema for b in blocks if b.num_columns > 0), None)
        if schema is not None:
            SortKey(key).validate_schema(schema)  # dataset.py:2172-2176: unknown key columns raise
    return GroupedData(blocks, key, num_partitions=num_partitions, data_context=data_context)

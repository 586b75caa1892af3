"""Block builders of the exchange path (host side).

Mirrors ``DelegatingBlockBuilder`` (python/ray/data/_internal/delegating_block_builder.py:
9-76), ``TableBlockBuilder`` / ``ArrowBlockBuilder`` (_internal/table_block.py:35-156,
_internal/arrow_block.py:130-160) and ``SizeEstimator`` (_internal/size_estimator.py:9-63).
The reference uses them in ``ShuffleTaskSpec.reduce`` (concatenate one slice per mapper)
and to rebuild aggregate outputs row by row; the engine's own reduce paths concatenate
device buffers instead, so these are the host-facing equivalents for code written against
the reference's builder API.
"""
from __future__ import annotations

import collections
import collections.abc
import pickle
from typing import Any, Dict, List, Mapping, Optional

import numpy as np
import pyarrow as pa

from .block import empty_table

# python/ray/data/_internal/table_block.py:25-28
MAX_UNCOMPACTED_SIZE_BYTES = 50 * 1024 * 1024


class _RunningMean:
    def __init__(self):
        self._weight = 0
        self._mean = 0.0

    def add(self, x: float, weight: int = 1) -> None:
        if weight == 0:
            return
        n1, n2 = self._weight, weight
        self._mean = (n1 * self._mean + n2 * x) / (n1 + n2)
        self._weight += weight

    @property
    def mean(self) -> float:
        return self._mean


class SizeEstimator:
    """Serialised-size estimate of a stream of items from a sample of them: the first 10,
    then every 10th up to 100, then every 100th (size_estimator.py:19-47).  The reference
    measures with Ray's serialiser; there is no object store here, so pickle stands in."""

    def __init__(self):
        self._running_mean = _RunningMean()
        self._count = 0

    def add(self, item: Any) -> None:
        self._count += 1
        if self._count <= 10:
            self._running_mean.add(self._real_size(item), weight=1)
        elif self._count <= 100:
            if self._count % 10 == 0:
                self._running_mean.add(self._real_size(item), weight=10)
        elif self._count % 100 == 0:
            self._running_mean.add(self._real_size(item), weight=100)

    def add_block(self, block: List[Any]) -> None:
        if self._count < 10:
            for i in range(min(10 - self._count, len(block))):
                self._running_mean.add(self._real_size(block[i]), weight=1)
        if self._count < 100:
            for i in range(10 - (self._count % 10), min(100 - self._count, len(block)), 10):
                self._running_mean.add(self._real_size(block[i]), weight=10)
        if (len(block) + (self._count % 100)) // 100 > 1:
            for i in range(100 - (self._count % 100), len(block), 100):
                self._running_mean.add(self._real_size(block[i]), weight=100)
        self._count += len(block)

    def size_bytes(self) -> int:
        return int(self._running_mean.mean * self._count)

    def _real_size(self, item: Any) -> int:
        return len(pickle.dumps(item, protocol=pickle.HIGHEST_PROTOCOL))


def _is_array_like(value: Any) -> bool:
    return hasattr(value, "__array__") or hasattr(value, "__array_interface__")


class ArrowBlockBuilder:
    """Rows (dicts) and whole tables in, one Arrow table out."""

    def __init__(self):
        self._columns: Dict[str, list] = collections.defaultdict(list)
        self._column_names = None
        self._tables: List[pa.Table] = []
        self._tables_size_cursor = 0
        self._tables_size_bytes = 0
        self._uncompacted_size = SizeEstimator()
        self._num_rows = 0
        self._num_compactions = 0

    def add(self, item: Mapping[str, Any]) -> None:
        if isinstance(item, np.ndarray):
            item = {"data": item}
        if not isinstance(item, collections.abc.Mapping):
            raise ValueError("Returned elements of an TableBlock must be of type `dict`, "
                             "got {} (type {}).".format(item, type(item)))
        item_column_names = item.keys()
        if self._column_names is not None:
            if item_column_names != self._column_names:
                raise ValueError(
                    "Current row has different columns compared to previous rows. "
                    f"Columns of current row: {sorted(item_column_names)}, "
                    f"Columns of previous rows: {sorted(self._column_names)}.")
        else:
            self._column_names = item_column_names
        for key, value in item.items():
            if _is_array_like(value) and not isinstance(value, np.ndarray):
                value = np.array(value)
            self._columns[key].append(value)
        self._num_rows += 1
        self._compact_if_needed()
        self._uncompacted_size.add(item)

    def add_block(self, block: pa.Table) -> None:
        if not isinstance(block, pa.Table):
            raise TypeError(f"Got a block of type {type(block)}, expected {pa.Table}."
                            "If you are mapping a function, ensure it returns an "
                            f"object with the expected type. Block:\n{block}")
        self._tables.append(block)
        self._num_rows += block.num_rows

    @staticmethod
    def _table_from_pydict(columns: Dict[str, List[Any]]) -> pa.Table:
        out = {}
        for name, values in columns.items():
            if values and isinstance(values[0], np.ndarray):
                out[name] = pa.array([v.tolist() for v in values])
            else:
                out[name] = pa.array(values)
        return pa.table(out)

    @staticmethod
    def _concat_tables(tables: List[pa.Table]) -> pa.Table:
        tables = [t for t in tables if t.num_columns > 0]
        if not tables:
            return empty_table()
        if len(tables) == 1:
            return tables[0]
        return pa.concat_tables(tables, promote_options="permissive")

    def will_build_yield_copy(self) -> bool:
        if self._columns:
            return True
        return False  # Arrow concatenation is zero-copy (arrow_block.py:_concat_would_copy)

    def build(self) -> pa.Table:
        tables = [self._table_from_pydict(self._columns)] if self._columns else []
        tables.extend(self._tables)
        if len(tables) > 0:
            return self._concat_tables(tables)
        return empty_table()

    def num_rows(self) -> int:
        return self._num_rows

    def get_estimated_memory_usage(self) -> int:
        if self._num_rows == 0:
            return 0
        for table in self._tables[self._tables_size_cursor:]:
            self._tables_size_bytes += table.nbytes
        self._tables_size_cursor = len(self._tables)
        return self._tables_size_bytes + self._uncompacted_size.size_bytes()

    def _compact_if_needed(self) -> None:
        assert self._columns
        if self._uncompacted_size.size_bytes() < MAX_UNCOMPACTED_SIZE_BYTES:
            return
        block = self._table_from_pydict(self._columns)
        self.add_block(block)
        self._num_rows -= block.num_rows  # add_block counted the rows a second time
        self._uncompacted_size = SizeEstimator()
        self._columns.clear()
        self._num_compactions += 1


class DelegatingBlockBuilder:
    """Picks the builder from the first thing added; an empty block is remembered and used
    only if nothing else arrives (delegating_block_builder.py:37-66)."""

    def __init__(self):
        self._builder: Optional[ArrowBlockBuilder] = None
        self._empty_block = None

    def add(self, item: Mapping[str, Any]) -> None:
        assert isinstance(item, collections.abc.Mapping), item
        if self._builder is None:
            self._builder = ArrowBlockBuilder()
        self._builder.add(item)

    def add_batch(self, batch) -> None:
        from .block_accessor import B200ArrowBlockAccessor

        if isinstance(batch, collections.abc.Mapping):
            batch = pa.table({k: pa.array(np.asarray(v).tolist() if np.asarray(v).ndim > 1 else np.asarray(v))
                              for k, v in batch.items()})
        self.add_block(B200ArrowBlockAccessor.for_block(batch).to_arrow())

    def add_block(self, block) -> None:
        from .block_accessor import B200ArrowBlockAccessor

        table = B200ArrowBlockAccessor.for_block(block).to_arrow()
        if table.num_rows == 0:
            self._empty_block = table
            return
        if self._builder is None:
            self._builder = ArrowBlockBuilder()
        self._builder.add_block(table)

    def will_build_yield_copy(self) -> bool:
        if self._builder is None:
            return True
        return self._builder.will_build_yield_copy()

    def build(self) -> pa.Table:
        if self._builder is None:
            self._builder = ArrowBlockBuilder()
            if self._empty_block is not None:
                self._builder.add_block(self._empty_block)
        return self._builder.build()

    def num_rows(self) -> int:
        return self._builder.num_rows() if self._builder is not None else 0

    def get_estimated_memory_usage(self) -> int:
        if self._builder is None:
            return 0
        return self._builder.get_estimated_memory_usage()

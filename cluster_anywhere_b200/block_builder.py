"""Block builder of the exchange path (host side).

The reference concatenates a reduce task's inputs through ``DelegatingBlockBuilder``
(python/ray/data/_internal/delegating_block_builder.py:9-76, used by
``ShuffleTaskSpec.reduce``, planner/exchange/shuffle_task_spec.py:115-118) and rebuilds
aggregate outputs row by row through ``ArrowBlockBuilder``
(_internal/table_block.py:35-156).  The engine needs far less: its outputs leave the
device as whole Arrow tables, so the builder here only has to put tables (and the odd
row dict from caller code written against the reference's builder API) together in
arrival order.  Same method names as the reference's builders; own implementation:
rows are buffered column-wise and turned into a table whenever a table arrives or the
block is built, sizes come from ``Table.nbytes`` (no sampling estimator, no object
store).
"""
from __future__ import annotations

from typing import Any, Dict, List, Mapping, Optional

import pyarrow as pa

from .block import empty_table


class BlockBuilder:
    def __init__(self):
        self._tables: List[pa.Table] = []   # finished pieces, arrival order
        self._cols: Optional[Dict[str, list]] = None  # pending rows, column-wise
        self._pending = 0
        self._rows = 0
        self._table_bytes = 0
        self._row_bytes = 0.0  # bytes per row of the pieces built from rows so far
        self._empty: Optional[pa.Table] = None  # schema donor when only empty blocks arrive

    # -- input -----------------------------------------------------------------
    def add(self, item: Any) -> None:
        """One row: a mapping column -> value (same columns for every row)."""
        if not isinstance(item, Mapping):
            raise ValueError(f"a row must be a mapping of column name to value, got {type(item).__name__}")
        if self._cols is None:
            self._cols = {str(k): [] for k in item}
        if item.keys() != self._cols.keys():
            raise ValueError(f"row has columns {sorted(item)}, the builder holds {sorted(self._cols)}")
        for k, v in item.items():
            self._cols[k].append(v)
        self._pending += 1
        self._rows += 1

    def add_block(self, block: Any) -> None:
        """One table (Arrow, or a pandas DataFrame, which is converted)."""
        if not isinstance(block, pa.Table):
            try:
                import pandas as pd
            except ImportError:  # pragma: no cover
                pd = None
            if pd is not None and isinstance(block, pd.DataFrame):
                block = pa.Table.from_pandas(block, preserve_index=False)
            else:
                raise TypeError(f"a block must be a pyarrow.Table or pandas.DataFrame, got {type(block).__name__}")
        if block.num_rows == 0:
            if self._empty is None and block.num_columns:
                self._empty = block
            return
        self._flush_rows()
        self._tables.append(block)
        self._rows += block.num_rows
        self._table_bytes += block.nbytes

    # -- output ----------------------------------------------------------------
    def _flush_rows(self) -> None:
        if self._pending:
            t = pa.table({k: pa.array(v) for k, v in self._cols.items()})
            self._row_bytes = t.nbytes / t.num_rows
            self._tables.append(t)
            self._table_bytes += t.nbytes
            for v in self._cols.values():
                v.clear()
            self._pending = 0

    def will_build_yield_copy(self) -> bool:
        """False only when ``build`` can hand back the single table it was given."""
        return not (len(self._tables) == 1 and self._pending == 0)

    def build(self) -> pa.Table:
        self._flush_rows()
        if not self._tables:
            return self._empty if self._empty is not None else empty_table()
        if len(self._tables) == 1:
            return self._tables[0]
        return pa.concat_tables(self._tables, promote_options="permissive")

    def num_rows(self) -> int:
        return self._rows

    def get_estimated_memory_usage(self) -> int:
        # pending rows: priced like the rows already converted (or 64 B before any were)
        return int(self._table_bytes + self._pending * (self._row_bytes or 64.0))


# the reference's names for the same role
ArrowBlockBuilder = BlockBuilder
DelegatingBlockBuilder = BlockBuilder

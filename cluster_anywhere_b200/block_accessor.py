"""``B200ArrowBlockAccessor``: the reference's Arrow block operations, on the GPU.

Mirrors the exchange-relevant part of ``ArrowBlockAccessor``
(python/ray/data/_internal/arrow_block.py:219-663) method for method -- same
names, arguments, return shapes and empty-block contract -- with every
arithmetic step (``sort_indices``, ``take``, ``searchsorted``, ``permutation``,
the per-group aggregate loops) replaced by sm_100a kernels behind the C-ABI.
Host Arrow tables go in, host Arrow tables come out (staging + kernels + copy
back per call); the device-resident, multi-block path is
``cluster_anywhere_b200.exchange.DeviceExchange``.
"""
from __future__ import annotations

import os

from typing import List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import torch

from . import device as dev
from . import keys as keymod
from .block import BlockExecStats, BlockMetadata, empty_table
from .sort_key import SortKey


def _device_sort(rt: dev.Runtime, dt: dev.DeviceTable, sort_key: SortKey, want_norm: bool):
    """-> (perm, norms); a string-run hint of the sort (dev.sort_permutation) rides on norms."""
    hints = {}
    perm, norms = dev.sort_permutation(rt, dt, sort_key.get_columns(), sort_key.get_descending(), want_norm, hints)
    if "string_runs" in hints:
        norms = _Norms(norms)
        norms.string_runs = hints["string_runs"]
    return perm, norms


class _Norms(list):
    """The radix-key list of a sort plus what else the sort learned about its key columns."""
    string_runs = None


def _presorted(sort_key: SortKey, norms):
    """Primary key column -> how dev.take_table can build it without a random gather: its
    radix keys in output order (numeric keys) or one source row per run of equal strings."""
    runs = getattr(norms, "string_runs", None)
    if runs is not None:
        return {runs[0]: ("rows", runs[1])}
    if not norms or norms[0] is None:
        return None
    return {sort_key.get_columns()[0]: (norms[0], sort_key.get_descending()[0])}


def _cuts_for(rt: dev.Runtime, dt: dev.DeviceTable, norms, boundaries, sort_key: SortKey,
              sorted_dt: Optional[dev.DeviceTable] = None) -> List[int]:
    """K4.  Numeric keys: search on the sorted radix keys.  String key columns are searched
    on the sorted column itself (``sorted_dt``), alone or inside a tuple with other columns."""
    cols, desc = sort_key.get_columns(), sort_key.get_descending()
    if any(dev.is_string_key(dt.column(c)) for c in cols):
        assert sorted_dt is not None
        if len(cols) != 1:
            return dev.mixed_partition_cuts(rt, sorted_dt, norms, boundaries, cols, desc)
        return dev.string_partition_cuts(rt, sorted_dt.column(cols[0]), boundaries, desc[0])
    ktypes = [dev.key_type_of(dt.column(c).type) for c in cols]
    lower, upper, flags = keymod.normalize_boundaries(boundaries, ktypes, desc)
    thr = [keymod.last_class_threshold(k, dt.column(c).null_count) for k, c in zip(ktypes, cols)]
    return dev.partition_cuts(rt, norms, dt.num_rows, lower, upper, flags, desc,
                              [t[0] for t in thr], [t[1] for t in thr])


class B200ArrowBlockAccessor:
    def __init__(self, table: pa.Table, runtime: Optional[dev.Runtime] = None):
        if not isinstance(table, pa.Table):
            raise TypeError(f"B200ArrowBlockAccessor needs a pyarrow.Table, got {type(table)}")
        self._table = table
        self._rt = runtime

    # -- plumbing ------------------------------------------------------------
    @staticmethod
    def for_block(block, runtime: Optional[dev.Runtime] = None) -> "B200ArrowBlockAccessor":
        """python/ray/data/block.py:446-473: Arrow tables and pandas DataFrames (a DataFrame
        gets the accessor whose outputs are DataFrames again)."""
        if isinstance(block, B200ArrowBlockAccessor):
            return block
        try:
            import pandas as pd

            if isinstance(block, pd.DataFrame):
                return B200PandasBlockAccessor(block, runtime)
        except ImportError:
            pass
        return B200ArrowBlockAccessor(block, runtime)

    @property
    def rt(self) -> dev.Runtime:
        if self._rt is None:
            self._rt = dev.default_runtime()
        return self._rt

    def to_arrow(self) -> pa.Table:
        return self._table

    def to_block(self) -> pa.Table:
        return self._table

    def num_rows(self) -> int:
        # arrow_block.py:286-289: a 0-column table counts as empty
        return self._table.num_rows if self._table.num_columns > 0 else 0

    def size_bytes(self) -> int:
        return self._table.nbytes

    def schema(self) -> pa.Schema:
        return self._table.schema

    def get_metadata(self, input_files=None, exec_stats: Optional[BlockExecStats] = None) -> BlockMetadata:
        return BlockMetadata(self.num_rows(), self.size_bytes(), self.schema(), input_files, exec_stats)

    @staticmethod
    def _empty_table() -> pa.Table:
        return empty_table()

    def slice(self, start: int, end: int, copy: bool = False) -> pa.Table:
        """arrow_block.py:219-223 (zero-copy view; ``copy`` compacts)."""
        view = self._table.slice(start, max(0, end - start))
        return view.combine_chunks() if copy else view

    def select(self, columns: List[str]) -> pa.Table:
        if not all(isinstance(c, str) for c in columns):
            raise ValueError(
                "Columns must be a list of column name strings when aggregating on "
                f"Arrow blocks, but got: {columns}."
            )
        return self._table.select(columns)

    # -- take / sample / shuffle ---------------------------------------------
    def take(self, indices) -> pa.Table:
        """Row gather on the GPU (arrow_block.py:319-328 -> transform_pyarrow.py:45-71)."""
        idx = np.ascontiguousarray(np.asarray(indices), dtype=np.int64)
        n = self.num_rows()
        if idx.size and (idx.min() < 0 or idx.max() >= n):
            raise IndexError("take index out of bounds")
        rt = self.rt
        dt = dev.stage_table(rt, self._table)
        idx_d = rt.empty(idx.size, torch.int32)
        dev._h2d(idx_d.view(torch.uint8), idx.astype(np.uint32).view(np.uint8))
        out = dev.take_table(rt, dt, idx_d, idx.size)
        return dev.export_table(rt, out)

    def sample(self, n_samples: int, sort_key: SortKey) -> pa.Table:
        """Key columns of ``n_samples`` random rows (table_block.py:253-264,
        arrow_block.py:341-344).  The reference samples with an unseeded
        ``random.sample``; so does this (the draw is host-side, the gather is K5)."""
        import random

        if sort_key is None or callable(sort_key):
            raise NotImplementedError(f"Table sort key must be a column name, was: {sort_key}")
        if self.num_rows() == 0:
            return self._empty_table()
        k = min(n_samples, self.num_rows())
        indices = random.sample(range(self.num_rows()), k)
        keys_only = B200ArrowBlockAccessor(self._table.select(sort_key.get_columns()), self._rt)
        return keys_only.take(indices)

    def random_shuffle(self, random_seed: Optional[int]) -> pa.Table:
        """arrow_block.py:225-231: a seeded random permutation of the rows.  The
        reference draws ``RandomState(seed).permutation(n)`` (MT19937, values unpinned
        by its tests); here the permutation is the Philox-keyed bijection of K7 and
        the gather is K5.  ``None`` draws a fresh seed."""
        n = self.num_rows()
        if n == 0:
            return self._table
        rt = self.rt
        dt = dev.stage_table(rt, self._table)
        return dev.export_table(rt, dev.shuffle_table(rt, dt, random_seed))

    # -- sort ------------------------------------------------------------------
    def sort(self, sort_key: SortKey) -> pa.Table:
        """arrow_block.py:425-438."""
        assert sort_key.get_columns(), (
            f"Sorting columns couldn't be empty (got {sort_key.get_columns()})"
        )
        if self._table.num_rows == 0:
            return self._empty_table()
        sort_key.validate_schema(self._table.schema)
        rt = self.rt
        dt = dev.stage_table(rt, self._table)
        perm, norms = _device_sort(rt, dt, sort_key, False)
        return dev.export_table(rt, dev.take_table(rt, dt, perm, dt.num_rows, _presorted(sort_key, norms)))

    def sort_and_partition(self, boundaries: List[tuple], sort_key: SortKey) -> List[pa.Table]:
        """arrow_block.py:440-450: sort, then ``len(boundaries)+1`` range partitions
        (util.py:806-828).  One staging, one sort, one gather, one copy back; the
        partitions are zero-copy slices of the sorted table, as in the reference."""
        assert sort_key.get_columns()
        if self._table.num_rows == 0:
            return [self._empty_table() for _ in range(len(boundaries) + 1)]
        sort_key.validate_schema(self._table.schema)
        rt = self.rt
        dt = dev.stage_table(rt, self._table)
        want = len(boundaries) > 0
        perm, norms = _device_sort(rt, dt, sort_key, want)
        sorted_dt = dev.take_table(rt, dt, perm, dt.num_rows, _presorted(sort_key, norms))
        if not want:
            return [dev.export_table(rt, sorted_dt)]
        cuts = _cuts_for(rt, dt, norms, boundaries, sort_key, sorted_dt)
        table = dev.export_table(rt, sorted_dt)
        out, last = [], 0
        for c in cuts:
            out.append(table[last:c])  # util.py:823-827 slices exactly like this
            last = c
        out.append(table[last:])
        return out

    @staticmethod
    def merge_sorted_blocks(blocks: List[pa.Table], sort_key: SortKey,
                            runtime: Optional[dev.Runtime] = None) -> Tuple[pa.Table, BlockMetadata]:
        """arrow_block.py:529-541 -> transform_pyarrow.py:540-547: drop empty runs,
        concatenate in mapper order (schemas unified with type promotion), stable
        re-sort.  Ties therefore resolve by (mapper index, position in run)."""
        stats = BlockExecStats.builder()
        blocks = [b for b in blocks if b.num_rows > 0]
        if len(blocks) == 0:
            ret = empty_table()
        else:
            t = pa.concat_tables(blocks, promote_options="permissive")
            acc = B200ArrowBlockAccessor(t, runtime)
            # the runs arrive sorted (SortTaskSpec.map made them): a stable merge (K6) gives
            # the reference's concat + stable re-sort; anything the merge kernel does not
            # take -- several key columns, string keys, 64-bit keys with nulls, a run that
            # turns out not to be sorted -- is sorted from scratch like the reference does
            ret = acc._merge_runs([b.num_rows for b in blocks], sort_key)
            if ret is None:
                ret = acc.sort(sort_key)
        return ret, B200ArrowBlockAccessor(ret, runtime).get_metadata(exec_stats=stats.build())

    def _merge_runs(self, run_rows: List[int], sort_key: SortKey) -> Optional[pa.Table]:
        """K6 on this table seen as consecutive sorted runs of ``run_rows`` rows; None when
        the merge kernel does not apply (see merge_sorted_blocks)."""
        import ctypes

        import torch

        cols, desc = sort_key.get_columns(), sort_key.get_descending()
        # a merge round costs 0.70 ms per 125 M keys, the full sort 4.1 ms
        # (profiles/r02_merge_vs_resort.jsonl): up to 32 runs (5 rounds) the merge wins
        if (len(cols) != 1 or len(run_rows) > int(os.environ.get("XCHG_MERGE_MAX_RUNS", "32"))
                or os.environ.get("XCHG_MERGE", "1") == "0"):
            return None
        sort_key.validate_schema(self._table.schema)
        rt = self.rt
        dt = dev.stage_table(rt, self._table)
        col = dt.column(cols[0])
        if col.kind != "fixed":
            return None
        try:
            kt = dev.key_type_of(col.type)
        except NotImplementedError:
            return None
        if col.validity is not None and kt in (dev._cabi.I64, dev._cabi.U64):
            return None  # nulls of 64-bit keys need the validity place of the full sort
        n = dt.num_rows
        if n > (1 << 30) - 1:
            return None
        norm = rt.empty(n, torch.int64)
        rt.call("xchg_normalize_keys", dev._vp(col.data), kt, int(bool(desc[0])), dev._vp(col.validity), 0, None, n,
                dev._vp(norm), rt.stream)
        starts = np.concatenate([[0], np.cumsum(np.asarray(run_rows, dtype=np.int64))])
        a_starts = (ctypes.c_uint64 * len(starts))(*[int(x) for x in starts])
        wsb = ctypes.c_uint64()
        dev._cabi.check(dev._cabi.load().xchg_merge_workspace_bytes(n, ctypes.byref(wsb)))
        ws = rt.empty(wsb.value)
        keys_out = rt.empty(n, torch.int64)
        perm = rt.empty(n, torch.int32)
        flag = rt.zeros(1, torch.int32)
        rt.call("xchg_merge_sorted_runs", dev._vp(norm), None, a_starts, len(run_rows), dev._vp(keys_out),
                dev._vp(perm), dev._vp(ws), wsb.value, dev._vp(flag), rt.stream,
                launches=max(1, (len(run_rows) - 1).bit_length()) + 1)
        if int(flag.item()):  # sync: a run was not sorted
            return None
        out = dev.take_table(rt, dt, perm, n, {cols[0]: (keys_out, bool(desc[0]))})
        return dev.export_table(rt, out)

    # -- aggregate -------------------------------------------------------------
    def _reduce(self, kind: int, on: str, ignore_nulls: bool, ddof: int = 1):
        """One global group over this block (the block reducers of arrow_block.py:346-423)."""
        from . import aggregate as agg

        cls = {agg.KIND_SUM: agg.Sum, agg.KIND_MIN: agg.Min, agg.KIND_MAX: agg.Max,
               agg.KIND_MEAN: agg.Mean}[kind]
        if self.num_rows() == 0:
            return None
        rt = self.rt
        dt = dev.stage_table(rt, self._table.select([on]))
        out = agg.aggregate_table(rt, dt, None, [cls(on, ignore_nulls=ignore_nulls)])
        return dev.export_table(rt, out).column(0)[0].as_py()

    def count(self, on: str) -> Optional[int]:
        """arrow_block.py:346-360: number of non-null values of ``on``."""
        if self.num_rows() == 0:
            return None
        col = self._table[on]
        return len(col) - col.null_count

    def sum(self, on: str, ignore_nulls: bool):
        from . import aggregate as agg

        return self._reduce(agg.KIND_SUM, on, ignore_nulls)

    def min(self, on: str, ignore_nulls: bool):
        from . import aggregate as agg

        return self._reduce(agg.KIND_MIN, on, ignore_nulls)

    def max(self, on: str, ignore_nulls: bool):
        from . import aggregate as agg

        return self._reduce(agg.KIND_MAX, on, ignore_nulls)

    def mean(self, on: str, ignore_nulls: bool):
        from . import aggregate as agg

        return self._reduce(agg.KIND_MEAN, on, ignore_nulls)

    def get_block_boundaries(self, keys) -> "np.ndarray":
        """Group boundaries of this block, which must be sorted by ``keys`` (a column name
        or a list): ``_get_block_boundaries`` (python/ray/data/block.py:515-592) as
        GroupedData.map_groups uses it (grouped_data.py:196-213)."""
        from . import aggregate as agg

        keys = agg.key_list(keys)
        if self.num_rows() == 0:
            return np.array([0, 0], dtype=int)
        rt = self.rt
        return agg.block_boundaries(rt, dev.stage_table(rt, self._table.select(keys)), keys)

    def unique(self, column: str) -> list:
        """Distinct values of ``column`` (``pac.unique`` of the reference's Unique aggregate,
        python/ray/data/aggregate.py:468-472), ascending, null last."""
        from . import aggregate as agg

        if self.num_rows() == 0:
            return []
        rt = self.rt
        dt = dev.stage_table(rt, self._table.select([column]))
        col = agg.unique_values(rt, dt, column)
        out = dev.export_table(rt, dev.DeviceTable([column], [col], col.n, pa.schema([(column, col.type)])))
        return out[column].to_pylist()

    @staticmethod
    def _group_key(sort_key: SortKey):
        """None (one global group), a column name, or the list of key columns."""
        cols = list(sort_key.get_columns()) if sort_key is not None else []
        if not cols:
            return None
        return cols[0] if len(cols) == 1 else cols

    def combine(self, sort_key: SortKey, aggs) -> pa.Table:
        """arrow_block.py:452-526: one row per group with the partially combined
        accumulators (K8).  The reference requires the block to be sorted by key and
        walks its runs; the hash aggregate does not need the order and emits the groups
        sorted by key, as the reference does.  ``sort_key`` without columns = one global
        group (arrow_block.py:473-476)."""
        from . import aggregate as agg

        key = self._group_key(sort_key)
        if self.num_rows() == 0:
            return self._empty_table()
        rt = self.rt
        dt = dev.stage_table(rt, agg.prune_columns(self._table, agg.key_list(key), aggs))
        out = agg.aggregate_table(rt, dt, key, aggs, finalize=False)
        return dev.export_table(rt, out)

    @staticmethod
    def aggregate_combined_blocks(blocks: List[pa.Table], sort_key: SortKey, aggs, finalize: bool = True,
                                  runtime: Optional[dev.Runtime] = None) -> Tuple[pa.Table, BlockMetadata]:
        """arrow_block.py:544-663: merge the partial blocks of one key range (K9) and
        finalize (``finalize=False`` keeps accumulators: the push-based merge stage,
        push_based_shuffle_task_scheduler.py:715)."""
        from . import aggregate as agg

        stats = BlockExecStats.builder()
        key = B200ArrowBlockAccessor._group_key(sort_key)
        blocks = [b for b in blocks if b.num_rows > 0 and b.num_columns > 0]
        if not blocks:
            ret = empty_table()
        else:
            rt = runtime or dev.default_runtime()
            t = pa.concat_tables(blocks, promote_options="default")
            out = agg.aggregate_table(rt, dev.stage_table(rt, t), key, aggs, finalize=finalize)
            ret = dev.export_table(rt, out)
            if finalize:
                ret = agg.null_typed_like_reference(ret)
        return ret, B200ArrowBlockAccessor(ret, runtime).get_metadata(exec_stats=stats.build())


def is_pandas_block(block) -> bool:
    try:
        import pandas as pd
    except ImportError:
        return False
    return isinstance(block, pd.DataFrame)


def want_pandas_output(blocks, batch_format: str = "default") -> bool:
    """Block type of a reduce task's output (TableBlockAccessor.normalize_block_types,
    python/ray/data/_internal/table_block.py:266-316): all inputs of one type -> that type;
    mixed -> ``batch_format`` decides."""
    kinds = {is_pandas_block(b) for b in blocks if b is not None}
    if len(kinds) == 1:
        return kinds.pop()
    return batch_format == "pandas"


def _to_pandas(table):
    import pandas as pd

    if isinstance(table, pd.DataFrame):
        return table
    if table.num_columns == 0:
        return pd.DataFrame()  # pandas_block.py: _empty_table()
    return table.to_pandas()


class B200PandasBlockAccessor(B200ArrowBlockAccessor):
    """Pandas twin (python/ray/data/_internal/pandas_block.py:503-746): DataFrame in,
    DataFrame out.  The rows go through the same kernels as Arrow blocks (the frame is
    converted on the way in -- missing values, NaN included, become nulls, which sort last
    in both directions like pandas' ``na_position='last'`` -- and back on the way out).
    The reference's single-key ``sort_values`` is not stable; this one is, which is one of
    the orders it may produce."""

    def __init__(self, frame, runtime: Optional[dev.Runtime] = None):
        super().__init__(pa.Table.from_pandas(frame, preserve_index=False), runtime)
        self._frame = frame

    def to_block(self):
        return self._frame

    def to_pandas(self):
        return self._frame

    @staticmethod
    def _empty_table():
        import pandas as pd

        return pd.DataFrame()

    def slice(self, start: int, end: int, copy: bool = False):
        # pandas_block.py:219-224: a re-indexed view, deep-copied on request
        view = self._frame.iloc[start:end].reset_index(drop=True)
        return view.copy(deep=True) if copy else view

    def select(self, columns: List[str]):
        return _to_pandas(super().select(columns))

    def take(self, indices):
        return _to_pandas(super().take(indices))

    def sample(self, n_samples: int, sort_key: SortKey):
        return _to_pandas(super().sample(n_samples, sort_key))

    def random_shuffle(self, random_seed: Optional[int]):
        return _to_pandas(super().random_shuffle(random_seed))

    def sort(self, sort_key: SortKey):
        return _to_pandas(super().sort(sort_key))

    def sort_and_partition(self, boundaries: List[tuple], sort_key: SortKey) -> list:
        return [_to_pandas(t) for t in super().sort_and_partition(boundaries, sort_key)]

    def combine(self, sort_key: SortKey, aggs):
        return _to_pandas(super().combine(sort_key, aggs))

    @staticmethod
    def merge_sorted_blocks(blocks: list, sort_key: SortKey, runtime: Optional[dev.Runtime] = None):
        tables = [B200ArrowBlockAccessor.for_block(b).to_arrow() for b in blocks]
        ret, meta = B200ArrowBlockAccessor.merge_sorted_blocks(tables, sort_key, runtime)
        return _to_pandas(ret), meta

    @staticmethod
    def aggregate_combined_blocks(blocks: list, sort_key: SortKey, aggs, finalize: bool = True,
                                  runtime: Optional[dev.Runtime] = None):
        tables = [B200ArrowBlockAccessor.for_block(b).to_arrow() for b in blocks]
        ret, meta = B200ArrowBlockAccessor.aggregate_combined_blocks(tables, sort_key, aggs, finalize, runtime)
        return _to_pandas(ret), meta

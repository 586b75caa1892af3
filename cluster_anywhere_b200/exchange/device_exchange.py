"""Device-resident exchange: the whole map -> exchange -> reduce of ``Dataset.sort``
on blocks that stay in HBM.

The reference runs M map tasks (sort + partition each block), ships M x R
partitions through the object store and runs R reduce tasks (concat + stable
re-sort) -- python/ray/data/_internal/planner/sort.py:29-79 driving
pull_based_shuffle_task_scheduler.py:84-120.  Its final output is fully
determined: the R output blocks are the range partitions (by the boundaries) of
the STABLE sort of the concatenation of the input blocks in (mapper index,
in-block position) order.  On one GPU that is computed
directly: blocks are staged back to back (so global row id order = (block, row)
order), one stable radix sort (K2+K3), one boundary search (K4), one payload
gather (K5).  With G GPUs each rank holds a contiguous run of blocks, sorts
locally, range-partitions by G-1 splitters and exchanges over NCCL (see
``sort`` below).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import torch

from .. import device as dev
from ..block import empty_table
from ..block_accessor import _cuts_for as cuts_for
from ..block_accessor import _presorted
from ..sort_key import SortKey
from .sort_task_spec import SortTaskSpec


@dataclass
class DeviceDataset:
    """Blocks resident in HBM: one concatenated device table plus the row range of
    every block (ranges follow Python slice semantics, as the reference's
    ``table[last:idx]`` partitions do)."""

    table: Optional[dev.DeviceTable]
    block_ranges: List[Tuple[int, int]]

    @property
    def num_rows(self) -> int:
        return 0 if self.table is None else self.table.num_rows

    @property
    def num_blocks(self) -> int:
        return len(self.block_ranges)


class DeviceExchange:
    def __init__(self, runtime: Optional[dev.Runtime] = None):
        self.rt = runtime or dev.default_runtime()

    # -- staging ---------------------------------------------------------------
    def stage(self, blocks: Sequence[pa.Table]) -> DeviceDataset:
        """Host Arrow blocks -> HBM (async copies on the current stream)."""
        nonempty = [b for b in blocks if b.num_rows > 0 and b.num_columns > 0]
        ranges, pos = [], 0
        for b in blocks:
            r = b.num_rows if b.num_columns > 0 else 0
            ranges.append((pos, pos + r))
            pos += r
        if not nonempty:
            return DeviceDataset(None, ranges)
        table = pa.concat_tables(nonempty, promote_options="permissive")  # zero-copy, chunked
        return DeviceDataset(dev.stage_table(self.rt, table), ranges)

    def export(self, ds: DeviceDataset) -> List[pa.Table]:
        """HBM -> host Arrow blocks (pinned host buffers, one per output block)."""
        if ds.table is None:
            return [empty_table() for _ in ds.block_ranges]
        # every block gets its own Arrow buffers (own int32 offsets: a block stays
        # under 2 GiB per column even when the device table does not); all copies
        # are queued first, then one stream synchronisation
        rt, t = self.rt, ds.table
        builders = []
        for lo, hi in ds.block_ranges:
            if hi <= lo:
                lo = hi = 0
            builders.append([dev._export_buffers(rt, c, lo, hi, []) for c in t.columns])
        rt.synchronize()
        return [pa.Table.from_arrays([b() for b in bs], schema=t.schema) for bs in builders]

    # -- sort --------------------------------------------------------------------
    def sample_boundaries(self, ds: DeviceDataset, sort_key: SortKey, num_reducers: int,
                          seed: int = 0) -> list:
        """K1 + the driver half of SortTaskSpec.sample_boundaries
        (sort_task_spec.py:154-226): ``int(R*10/M)`` rows per block, key columns
        gathered on the device, quantiles picked on the host.  The reference draws
        with an unseeded ``random.sample``; the draw here is Philox-seeded so that
        runs are reproducible (cut points are unpinned by the reference)."""
        if ds.table is None or ds.num_rows == 0:
            return [None] * (num_reducers - 1)
        m = max(1, ds.num_blocks)
        n_samples = int(num_reducers * 10 / m)
        picks = []
        for b, (lo, hi) in enumerate(ds.block_ranges):
            rows = hi - lo
            k = min(n_samples, rows)
            if k <= 0:
                continue
            g = np.random.Generator(np.random.Philox(key=[seed, b]))
            picks.append(lo + g.choice(rows, size=k, replace=False))
        if not picks:
            return [None] * (num_reducers - 1)
        idx = np.concatenate(picks).astype(np.uint32)
        rt = self.rt
        idx_d = rt.empty(idx.size, torch.int32)
        dev._h2d(idx_d.view(torch.uint8), idx.view(np.uint8))
        cols = sort_key.get_columns()
        key_cols = [dev.take_column(rt, ds.table.column(c), idx_d, idx.size) for c in cols]
        kt = dev.DeviceTable(list(cols), key_cols, idx.size,
                             pa.schema([ds.table.schema.field(c) for c in cols]))
        host = dev.export_table(rt, kt)
        tuples = list(zip(*[host[c].to_pylist() for c in cols]))
        return SortTaskSpec.boundaries_from_samples(tuples, sort_key, num_reducers)

    def sort(self, ds: DeviceDataset, sort_key: SortKey, num_outputs: Optional[int] = None,
             boundaries: Optional[list] = None, seed: int = 0) -> DeviceDataset:
        """``Dataset.sort`` on resident blocks -> resident, range-partitioned output.

        ``num_outputs`` defaults to the number of input blocks
        (python/ray/data/_internal/planner/sort.py:43-45); explicit ``boundaries``
        are boundary tuples in sort order (sort.py:55-62 builds them from
        ``SortKey.boundaries``)."""
        assert sort_key.get_columns()
        r = num_outputs if num_outputs is not None else max(1, ds.num_blocks)
        if boundaries is not None:
            r = len(boundaries) + 1
        if ds.table is None or ds.num_rows == 0:
            return DeviceDataset(None, [(0, 0)] * r)
        sort_key.validate_schema(ds.table.schema)
        rt = self.rt
        if boundaries is None:
            boundaries = self.sample_boundaries(ds, sort_key, r, seed) if r > 1 else []
        n = ds.num_rows
        perm, norms = dev.sort_permutation(rt, ds.table, sort_key.get_columns(),
                                           sort_key.get_descending(), want_norm=len(boundaries) > 0)
        cuts = cuts_for(rt, ds.table, norms, boundaries, sort_key) if boundaries else []
        out = dev.take_table(rt, ds.table, perm, n, _presorted(sort_key, norms))
        ranges, last = [], 0
        for c in cuts:
            ranges.append((last, c))
            last = c
        ranges.append((last, n))
        return DeviceDataset(out, ranges)

    def sort_blocks(self, blocks: Sequence[pa.Table], sort_key: SortKey,
                    num_outputs: Optional[int] = None, boundaries: Optional[list] = None,
                    seed: int = 0) -> List[pa.Table]:
        """Host blocks in, sorted host blocks out (stage + sort + export)."""
        return self.export(self.sort(self.stage(blocks), sort_key, num_outputs, boundaries, seed))

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

import os
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
    def __init__(self, runtime: Optional[dev.Runtime] = None, trace: bool = False):
        self.rt = runtime or dev.default_runtime()
        # trace=True: synchronise after every phase and accumulate wall-clock ms per
        # phase in ``phase_ms`` (diagnostics only -- it serialises the pipeline)
        self.trace = trace
        self.phase_ms = {}
        self._t_last = None
        # XCHG_TIMELINE=1: host wall-clock between stamps WITHOUT synchronising (shows
        # where the host thread blocks: collectives, .tolist() / .cpu() read-backs)
        self.timeline = os.environ.get("XCHG_TIMELINE", "0") == "1"
        self.timeline_ms = {}
        self._tl_last = None

    def _gpu_mark(self, name: str):
        """XCHG_TIMELINE=1: record a CUDA event on the current stream (no sync); bench.py
        turns the events of a step into offsets from the step's first mark."""
        if not self.timeline or not self.rt.is_cuda:
            return
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(self.rt.device))
        if not hasattr(self, "gpu_marks"):
            self.gpu_marks = []
        self.gpu_marks.append((name, ev))

    def gpu_timeline(self) -> dict:
        """Mean offset (ms) of every mark from the preceding 'step_begin' mark."""
        marks = getattr(self, "gpu_marks", [])
        if not marks:
            return {}
        torch.cuda.synchronize()
        acc, cnt, t0 = {}, {}, None
        for name, ev in marks:
            if name == "step_begin":
                t0 = ev
                continue
            if t0 is None:
                continue
            acc[name] = acc.get(name, 0.0) + t0.elapsed_time(ev)
            cnt[name] = cnt.get(name, 0) + 1
        return {k: round(acc[k] / cnt[k], 2) for k in acc}

    def _stamp(self, name: Optional[str]):
        if not self.timeline:
            return
        import time

        now = time.perf_counter()
        if name is not None and self._tl_last is not None:
            self.timeline_ms[name] = self.timeline_ms.get(name, 0.0) + (now - self._tl_last) * 1e3
        self._tl_last = now

    def _take_and_exchange(self, table, perm, send_counts, group):
        """K5 + exchange.  On a GPU the gather kernels store straight into the
        destination ranks' receive arenas over NVLink (exchange/peer.py); otherwise (host
        double under gloo, XCHG_P2P=0, or no symmetric memory on this platform) the rows
        are gathered locally and cross in one all-to-all-v per buffer."""
        from . import collective as coll
        from . import peer

        rt = self.rt
        if peer.enabled(rt, group) and not getattr(self, "_p2p_unavailable", False):
            try:
                recv = peer.gather_exchange(rt, self, table, perm, send_counts, group)
                self._phase("dist.gather_exchange_p2p")
                return recv
            except (ImportError, AttributeError, NotImplementedError) as e:  # no symmetric memory here
                import warnings

                warnings.warn(f"NVLink peer-memory exchange unavailable ({e!r}); using NCCL all-to-all")
                self._p2p_unavailable = True
        n = table.num_rows
        send_table = dev.take_table(rt, table, perm, n) if n > 0 else table
        self._phase("dist.gather_send")
        recv = coll.exchange_table(rt, send_table, send_counts, group)
        self._phase("dist.all_to_all")
        return recv

    def _phase(self, name: Optional[str]):
        if not self.trace:
            return
        import time

        self.rt.synchronize()
        now = time.perf_counter()
        if name is not None and self._t_last is not None:
            self.phase_ms[name] = self.phase_ms.get(name, 0.0) + (now - self._t_last) * 1e3
        self._t_last = now

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
        # one device->host copy per column buffer; blocks are windows into it
        return dev.export_blocks(self.rt, ds.table, ds.block_ranges)

    # -- sort --------------------------------------------------------------------
    def sample_keys(self, ds: DeviceDataset, sort_key: SortKey, n_samples: int, seed: int = 0,
                    first_block: int = 0) -> List[tuple]:
        """K1: key tuples of ``min(n_samples, rows)`` random rows of every block
        (SortTaskSpec._sample_block, sort_task_spec.py:229 ->
        ArrowBlockAccessor._sample, arrow_block.py:341-344).  Row indices are drawn on
        the host (Philox, keyed by (seed, global block index)); the key columns are
        gathered on the device (K5) and only the samples come back."""
        if ds.table is None or ds.num_rows == 0 or n_samples <= 0:
            return []
        picks = []
        for b, (lo, hi) in enumerate(ds.block_ranges):
            rows = hi - lo
            k = min(n_samples, rows)
            if k <= 0:
                continue
            g = np.random.Generator(np.random.Philox(key=[seed, first_block + b]))
            picks.append(lo + g.choice(rows, size=k, replace=False))
        if not picks:
            return []
        idx = np.concatenate(picks).astype(np.uint32)
        rt = self.rt
        idx_d = rt.empty(idx.size, torch.int32)
        dev._h2d(idx_d.view(torch.uint8), idx.view(np.uint8))
        cols = sort_key.get_columns()
        key_cols = [dev.take_column(rt, ds.table.column(c), idx_d, idx.size) for c in cols]
        kt = dev.DeviceTable(list(cols), key_cols, idx.size,
                             pa.schema([ds.table.schema.field(c) for c in cols]))
        host = dev.export_table(rt, kt)
        return list(zip(*[host[c].to_pylist() for c in cols]))

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
        tuples = self.sample_keys(ds, sort_key, int(num_reducers * 10 / m), seed)
        return SortTaskSpec.boundaries_from_samples(tuples, sort_key, num_reducers)

    def sort(self, ds: DeviceDataset, sort_key: SortKey, num_outputs: Optional[int] = None,
             boundaries: Optional[list] = None, seed: int = 0, _complete=None) -> DeviceDataset:
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
        self._stamp("tl.sort_prologue")
        hints = {}
        perm, norms = dev.sort_permutation(rt, ds.table, sort_key.get_columns(),
                                           sort_key.get_descending(), want_norm=len(boundaries) > 0, hints=hints)
        if "string_runs" in hints:
            from ..block_accessor import _Norms

            norms = _Norms(norms)
            norms.string_runs = hints["string_runs"]
        self._stamp("tl.sort_permutation")
        self._gpu_mark("g.keys_sorted")
        string_key = any(dev.is_string_key(ds.table.column(c)) for c in sort_key.get_columns())
        cuts = cuts_for(rt, ds.table, norms, boundaries, sort_key) if (boundaries and not string_key) else []
        self._stamp("tl.cuts")
        table = ds.table if _complete is None else _complete()  # payload columns may arrive late
        self._stamp("tl.complete")
        out = dev.take_table(rt, table, perm, n, _presorted(sort_key, norms))
        self._stamp("tl.take_launch")
        self._gpu_mark("g.taken")
        if boundaries and string_key:  # string keys are searched on the sorted column itself
            cuts = cuts_for(rt, ds.table, norms, boundaries, sort_key, out)
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
        self._phase(None)
        ds = self.stage(blocks)
        self._phase("e2e.stage_h2d")
        out = self.sort(ds, sort_key, num_outputs, boundaries, seed)
        self._phase("e2e.sort")
        res = self.export(out)
        self._phase("e2e.export_d2h")
        return res

    # -- multi-GPU sort ------------------------------------------------------------
    def sort_distributed(self, ds: DeviceDataset, sort_key: SortKey,
                         boundaries: Optional[list] = None, seed: int = 0,
                         group=None, force_exchange: bool = False) -> DeviceDataset:
        """``Dataset.sort`` over the blocks of ALL ranks (one process per GPU).

        Rank g holds a contiguous run of the dataset's input blocks (rank order =
        block order).  The global result is what the reference computes: the R range
        partitions, by ``boundaries``, of the stable sort of all rows in (block, row)
        order; rank g ends up with a contiguous run of the R output blocks.

          map     (SortTaskSpec.map, sort_task_spec.py:127-138) every row gets its
                  destination rank from the G-1 splitters -- the block boundaries at
                  the rank cuts -- with the reference's own partition rule stated per
                  row (xchg_assign_partitions); a stable one-digit radix pass orders
                  the row ids by destination; K5 gathers every column in that order.
          exchange  one all-to-all-v per column buffer over NVLink instead of M x R
                  plasma objects (pull_based_shuffle_task_scheduler.py:84-120).
          reduce  (SortTaskSpec.reduce, :140-152 -> merge_sorted_blocks) rows arrive in
                  (source rank, source position) order; one stable sort (K2+K3), the
                  boundary search for this rank's block cuts (K4), one gather (K5).
        """
        from . import collective as coll
        from . import chunked_sort, peer
        from .comm import as_comm

        group = as_comm(group)
        rank, G = group.rank, group.world
        if G == 1 and not force_exchange:  # force_exchange: run the map/exchange/reduce path anyway
            return self.sort(ds, sort_key, boundaries=boundaries, seed=seed)
        rt = self.rt
        self._phase(None)
        self._stamp(None)
        # chunked fast path (one fixed-width key column, fixed / uniform-width payload); the
        # ranks decide together, a None means "not eligible somewhere": general path below
        if (peer.enabled(rt, group) and not self.trace and not getattr(self, "_p2p_unavailable", False)
                and os.environ.get("XCHG_CHUNKED", "1") != "0"):
            try:
                out = chunked_sort.try_sort(self, ds, sort_key, boundaries, seed, group)
            except (ImportError, AttributeError, NotImplementedError) as e:  # no symmetric memory here
                import warnings

                warnings.warn(f"NVLink peer-memory exchange unavailable ({e!r}); using NCCL all-to-all")
                self._p2p_unavailable = True
                out = None
            if out is not None:
                return out
        self._gpu_mark("step_begin")
        schema = None if ds.table is None else ds.table.schema
        info = coll.gather_dataset_info(rt, ds.num_blocks, ds.num_rows, schema, group)
        self._gpu_mark("g.info_gathered")
        self._phase("dist.gather_info")
        self._stamp("tl.gather_info")
        blocks_per_rank = [i[0] for i in info]
        total_rows = sum(i[1] for i in info)
        schema = next((i[2] for i in info if i[2] is not None), None)
        r_total = sum(blocks_per_rank) if boundaries is None else len(boundaries) + 1
        r_total = max(1, r_total)
        # output blocks per rank: as many as it brought in; evenly when that would
        # leave a gap (user boundaries, or a rank without input blocks)
        if boundaries is None and all(b > 0 for b in blocks_per_rank):
            out_per_rank = list(blocks_per_rank)
        else:
            out_per_rank = [r_total // G + (1 if g < r_total % G else 0) for g in range(G)]
        ob = np.concatenate([[0], np.cumsum(out_per_rank)]).astype(int)
        my_blocks = int(out_per_rank[rank])
        if total_rows == 0 or schema is None:
            return DeviceDataset(None, [(0, 0)] * my_blocks)
        sort_key.validate_schema(schema)
        table = ds.table if ds.table is not None else coll.empty_device_table(rt, schema)

        if boundaries is None:
            first_block = int(sum(blocks_per_rank[:rank]))
            m = max(1, sum(blocks_per_rank))
            mine = self.sample_keys(ds, sort_key, int(r_total * 10 / m), seed, first_block)
            self._stamp("tl.sample_keys")
            samples = [t for part in coll.all_gather_object(mine, group) for t in part]
            self._stamp("tl.gather_samples")
            self._gpu_mark("g.samples_gathered")
            boundaries = (SortTaskSpec.boundaries_from_samples(samples, sort_key, r_total)
                          if r_total > 1 else [])
        self._phase("dist.sample_boundaries")
        ncols = len(sort_key.get_columns())
        none_tuple = (None,) * ncols
        # splitters sit between consecutive ranks that own output blocks: the block
        # boundary at the end of the earlier owner's run
        owners = [g for g in range(G) if out_per_rank[g] > 0]
        splitters = [boundaries[int(ob[g + 1]) - 1] for g in owners[:-1]]
        splitters = [none_tuple if s is None else s for s in splitters]

        # ---- map: destination of every row, stable order by destination ----
        n = table.num_rows
        send_counts = [0] * G
        perm = None
        if n > 0:
            dest, counts = dev.assign_partitions(rt, table, splitters, sort_key)
            perm = dev.sort_u32(rt, dest, n)
            for d, cnt in enumerate(counts.tolist()):
                send_counts[owners[d]] = int(cnt)
            self._phase("dist.assign_and_order")
            self._stamp("tl.assign_and_order")
            self._gpu_mark("g.assigned")
        lo_b, hi_b = int(ob[rank]), int(ob[rank + 1])
        local_bounds = list(boundaries[lo_b:max(lo_b, hi_b - 1)])
        if (peer.enabled(rt, group) and not getattr(self, "_p2p_unavailable", False) and not self.trace
                and os.environ.get("XCHG_OVERLAP", "1") != "0" and rt.is_cuda):
            try:
                return self._exchange_and_reduce_overlapped(table, perm, send_counts, sort_key,
                                                            local_bounds, my_blocks, group)
            except (ImportError, AttributeError, NotImplementedError) as e:
                import warnings

                warnings.warn(f"NVLink peer-memory exchange unavailable ({e!r}); using NCCL all-to-all")
                self._p2p_unavailable = True
        recv = self._take_and_exchange(table, perm, send_counts, group)

        # ---- reduce: stable sort of the received runs, cut into this rank's blocks ----
        if my_blocks == 0:
            assert recv.num_rows == 0
            return DeviceDataset(None, [])
        out = self.sort(DeviceDataset(recv, [(0, recv.num_rows)]), sort_key, boundaries=local_bounds)
        self._phase("dist.reduce_sort")
        return out

    def _exchange_and_reduce_overlapped(self, table, perm, send_counts, sort_key, local_bounds,
                                        my_blocks, group) -> DeviceDataset:
        """Exchange + reduce with the payload crossing NVLink WHILE the received keys are
        sorted (the push-based schedule's idea -- merge work pipelined with transfers,
        push_based_shuffle_task_scheduler.py:548-561 -- at the granularity this machine
        offers: two streams).

          main stream   key columns -> peers | fence | K2 + K3 on the received keys, K4 cuts
          side stream                 payload columns -> peers (few CTAs per SM: the
                                      stores are link-bound) | fence
          main stream   waits for the side stream, then the final K5 over all columns
        """
        from . import peer

        rt = self.rt
        pl = peer.plan_exchange(rt, self, table, perm, send_counts, group)
        self._stamp("tl.plan_exchange")
        names = list(table.names)
        key_idx = [names.index(c) for c in sort_key.get_columns()]
        rest_idx = [j for j in range(len(names)) if j not in key_idx]
        main = torch.cuda.current_stream(rt.device)
        if getattr(self, "_side_stream", None) is None:
            # XCHG_SIDE_PRIORITY=1 gives the payload stores a high-priority stream; measured
            # at 8 GPUs it only moves the wait from the stores to the key sort (42.4 vs
            # 41.0 ms per step), so the default stays at equal priority
            prio = -1 if os.environ.get("XCHG_SIDE_PRIORITY", "0") == "1" else 0
            self._side_stream = torch.cuda.Stream(device=rt.device, priority=prio)
        side = self._side_stream

        self._gpu_mark("g.planned")

        def payload_stores():
            side.wait_stream(main)
            with torch.cuda.stream(side):
                rt.call("xchg_set_gather_ctas_per_sm", int(os.environ.get("XCHG_SIDE_CTAS_PER_SM", "2")), launches=0)
                try:
                    peer.store_columns(rt, pl, rest_idx)
                finally:
                    rt.call("xchg_set_gather_ctas_per_sm", 0, launches=0)
                self._gpu_mark("g.side_payload_stored")

        def payload_fence():
            with torch.cuda.stream(side):
                peer.fence(rt, group)
                self._gpu_mark("g.side_payload_fenced")

        # Keys first, then the payload (92 % of the bytes) on the side stream while the
        # received keys are sorted.  XCHG_PAYLOAD_FIRST=1 queues the payload stores before
        # the key columns instead (measured at 8 GPUs: 43.8 vs 41.0 ms per step -- the keys
        # then share the links and the sort starts later).  The two fences are collectives
        # of one communicator and complete in issue order, so the key fence is always issued
        # BEFORE the payload fence (else the key sort would wait for the payload).
        payload_first = os.environ.get("XCHG_PAYLOAD_FIRST", "0") == "1"
        if payload_first:
            payload_stores()
        peer.store_columns(rt, pl, key_idx)
        self._gpu_mark("g.keys_stored")
        peer.fence(rt, group)
        self._gpu_mark("g.keys_fenced")
        if not payload_first:
            payload_stores()
        payload_fence()
        self._stamp("tl.launch_stores")

        m = int(pl.recv_rows[pl.rank])
        if my_blocks == 0 or m == 0:
            main.wait_stream(side)
            return DeviceDataset(None, [(0, 0)] * my_blocks)
        key_cols = peer.received_columns(rt, pl, key_idx)
        keys_table = dev.DeviceTable([names[j] for j in key_idx], key_cols, m,
                                     pa.schema([table.schema.field(j) for j in key_idx]))

        def complete():
            main.wait_stream(side)
            rest_cols = peer.received_columns(rt, pl, rest_idx)
            by_idx = dict(zip(key_idx, key_cols))
            by_idx.update(zip(rest_idx, rest_cols))
            return dev.DeviceTable(names, [by_idx[j] for j in range(len(names))], m, table.schema)

        out = self.sort(DeviceDataset(keys_table, [(0, m)]), sort_key, boundaries=local_bounds,
                        _complete=complete)
        self._stamp("tl.reduce_sort_launch")
        return out

    def sort_blocks_distributed(self, blocks: Sequence[pa.Table], sort_key: SortKey,
                                boundaries: Optional[list] = None, seed: int = 0,
                                group=None) -> List[pa.Table]:
        """This rank's host blocks in, this rank's run of the sorted output blocks out."""
        return self.export(self.sort_distributed(self.stage(blocks), sort_key, boundaries, seed, group))

    # -- random shuffle --------------------------------------------------------------
    def random_shuffle(self, ds: DeviceDataset, seed: Optional[int],
                       num_outputs: Optional[int] = None, group=None,
                       force_exchange: bool = False) -> DeviceDataset:
        """``Dataset.random_shuffle(seed)`` on resident blocks (all ranks' blocks when a
        process group is up).

        The reference permutes every block, deals its rows into R equal slices, sends
        slice j to reducer j and permutes again there (shuffle_task_spec.py:81-123):
        a random, even spread of every block over the R outputs followed by a random
        order inside each output.  Here ONE seeded permutation P of all N rows (K7)
        gives row g (global id = rows of earlier ranks + local row) the output
        position P(g); output block j is the even slice [j*N/R, (j+1)*N/R) of that
        order, R = number of input blocks (random_shuffle.py:29-83).  Same seed ->
        same output, whatever the number of GPUs.

        One GPU: ``take(P^-1(0..N))``.  G GPUs: every rank evaluates P on its own rows,
        orders them by destination block (stable one/two-digit radix pass), gathers
        (K5), all-to-all-v over NVLink, and the receiver inverts the received
        positions into its gather index."""
        from . import collective as coll

        rt = self.rt
        rank, G = coll.world(group)
        seed64 = dev.resolve_seed(seed)
        if G > 1 and seed is None:  # every rank must use the same permutation
            seed64 = coll.all_gather_object(seed64, group)[0]
        schema = None if ds.table is None else ds.table.schema
        if G == 1 and not force_exchange:
            info = [(ds.num_blocks, ds.num_rows, schema)]
        else:
            info = coll.gather_dataset_info(rt, ds.num_blocks, ds.num_rows, schema, group)
        n_total = sum(i[1] for i in info)
        r_total = max(1, num_outputs if num_outputs is not None else sum(i[0] for i in info))
        schema = next((i[2] for i in info if i[2] is not None), None)
        blocks_per_rank = [i[0] for i in info]
        if num_outputs is None and all(b > 0 for b in blocks_per_rank):
            out_per_rank = list(blocks_per_rank)
        else:
            out_per_rank = [r_total // G + (1 if g < r_total % G else 0) for g in range(G)]
        ob = np.concatenate([[0], np.cumsum(out_per_rank)]).astype(int)
        my_blocks = int(out_per_rank[rank])
        if n_total == 0 or schema is None:
            return DeviceDataset(None, [(0, 0)] * my_blocks)
        starts = [(j * n_total) // r_total for j in range(r_total + 1)]  # block j = [starts[j], starts[j+1])
        base = starts[int(ob[rank])]
        my_ranges = [(starts[j] - base, starts[j + 1] - base) for j in range(int(ob[rank]), int(ob[rank + 1]))]

        if G == 1 and not force_exchange:
            idx = dev.shuffle_positions(rt, seed64, n_total, 0, n_total, inverse=True)
            return DeviceDataset(dev.take_table(rt, ds.table, idx, n_total), my_ranges)

        table = ds.table if ds.table is not None else coll.empty_device_table(rt, schema)
        n = table.num_rows
        first = int(sum(i[1] for i in info[:rank]))
        send_counts = [0] * G
        pos_col = dev.DeviceColumn(pa.uint32(), n, "fixed", 4, rt.empty(0))
        perm = None
        if n > 0:
            pos = dev.shuffle_positions(rt, seed64, n_total, first, n, inverse=False)
            dest = rt.empty(n, torch.int32)
            counts = rt.zeros(r_total, torch.int64)
            rt.call("xchg_even_split_dest", dev._vp(pos), n, n_total, r_total, dev._vp(dest),
                    dev._vp(counts), rt.stream)
            perm = dev.sort_u32(rt, dest, n)
            per_block = counts.tolist()
            for g in range(G):
                send_counts[g] = int(sum(per_block[int(ob[g]):int(ob[g + 1])]))
            pos_col = dev.DeviceColumn(pa.uint32(), n, "fixed", 4, pos.view(torch.uint8))
        # the output position travels with its row as one more fixed-width column
        wire = dev.DeviceTable(list(table.names) + ["__xchg_pos"], list(table.columns) + [pos_col],
                               n, table.schema.append(pa.field("__xchg_pos", pa.uint32())))
        recv = self._take_and_exchange(wire, perm, send_counts, group)
        m = recv.num_rows
        if my_blocks == 0 or m == 0:
            assert m == 0
            return DeviceDataset(None, [(0, 0)] * my_blocks)
        inv = rt.empty(m, torch.int32)
        rt.call("xchg_invert_permutation", dev._vp(recv.columns[-1].data), m, base, dev._vp(inv), rt.stream)
        body = dev.DeviceTable(list(table.names), recv.columns[:-1], m, table.schema)
        return DeviceDataset(dev.take_table(rt, body, inv, m), my_ranges)

    def random_shuffle_blocks(self, blocks: Sequence[pa.Table], seed: Optional[int],
                              num_outputs: Optional[int] = None, group=None) -> List[pa.Table]:
        """Host blocks in, shuffled host blocks out (this rank's run when distributed)."""
        return self.export(self.random_shuffle(self.stage(blocks), seed, num_outputs, group))

    # -- group-by aggregate ------------------------------------------------------------
    def aggregate(self, ds: DeviceDataset, key: Optional[str], aggs: Sequence,
                  boundaries: Optional[list] = None, seed: int = 0, group=None,
                  force_exchange: bool = False) -> DeviceDataset:
        """``Dataset.groupby(key).aggregate(*aggs)`` on resident blocks (all ranks'
        blocks when a process group is up).

        Reference: generate_aggregate_fn (python/ray/data/_internal/planner/aggregate.py:
        42-98): boundaries sampled on the key, R = number of input blocks (``key=None``:
        one output block with one row, :61-63), then SortAggregateTaskSpec map / reduce
        under the pull- or push-based schedule.  The result is fully determined: one row
        per distinct key, sorted by key, cut into R range partitions by the boundaries.

        One GPU: ONE group aggregate over all resident rows (K8: direct-address, hash or
        sorted grouping by key type) and the boundary search
        over the sorted groups (K4).  G GPUs: local aggregate -> partial rows (one
        per local group) -> range split by the G-1 splitters (K4) -> all-to-all-v of the
        partials over NVLink (a few bytes per GROUP, not per row) -> merge (K9) ->
        finalize -> this rank's run of output blocks."""
        from .. import aggregate as agg
        from . import collective as coll

        rt = self.rt
        rank, G = coll.world(group)
        dist_path = G > 1 or force_exchange
        agg.check_aggs(aggs)
        schema = None if ds.table is None else ds.table.schema
        info = (coll.gather_dataset_info(rt, ds.num_blocks, ds.num_rows, schema, group) if dist_path
                else [(ds.num_blocks, ds.num_rows, schema)])
        blocks_per_rank = [i[0] for i in info]
        total_rows = sum(i[1] for i in info)
        schema = next((i[2] for i in info if i[2] is not None), None)
        if key is None:
            r_total = 1
        else:
            r_total = max(1, sum(blocks_per_rank) if boundaries is None else len(boundaries) + 1)
        if boundaries is None and key is not None and all(b > 0 for b in blocks_per_rank):
            out_per_rank = list(blocks_per_rank)
        else:
            out_per_rank = [r_total // G + (1 if g < r_total % G else 0) for g in range(G)]
        ob = np.concatenate([[0], np.cumsum(out_per_rank)]).astype(int)
        my_blocks = int(out_per_rank[rank])
        if total_rows == 0 or schema is None:
            return DeviceDataset(None, [(0, 0)] * my_blocks)
        for a in aggs:
            a._validate(schema)
        sk = SortKey(key) if key is not None else SortKey()
        if key is not None:
            sk.validate_schema(schema)
            if boundaries is None:
                first_block = int(sum(blocks_per_rank[:rank]))
                m = max(1, sum(blocks_per_rank))
                mine = self.sample_keys(ds, sk, int(r_total * 10 / m), seed, first_block)
                samples = ([t for part in coll.all_gather_object(mine, group) for t in part]
                           if dist_path else mine)
                boundaries = (SortTaskSpec.boundaries_from_samples(samples, sk, r_total)
                              if r_total > 1 else [])
            nkeys = len(sk.get_columns())
            boundaries = [(None,) * nkeys if b is None else b for b in boundaries]
        else:
            boundaries = []

        def cut_ranges(result, bounds, nblocks):
            if result is None:
                return [(0, 0)] * nblocks
            cuts = agg.group_cut_points(rt, result, key, bounds) if (key is not None and bounds) else []
            ranges, last = [], 0
            for c in cuts:
                ranges.append((last, c))
                last = c
            ranges.append((last, result.num_rows))
            return ranges

        if not dist_path:
            out = agg.aggregate_table(rt, ds.table, key, aggs, finalize=True)
            return DeviceDataset(out, cut_ranges(out, boundaries, r_total))

        # ---- map: local partial aggregate, split by destination rank ----
        owners = [g for g in range(G) if out_per_rank[g] > 0]
        partial = (agg.aggregate_table(rt, ds.table, key, aggs, finalize=False)
                   if ds.table is not None and ds.num_rows > 0 else None)
        pschemas = coll.all_gather_object(None if partial is None else partial.schema, group)
        pschema = next(s for s in pschemas if s is not None)
        send_counts = [0] * G
        if partial is None:
            partial = coll.empty_device_table(rt, pschema)
        elif key is None:
            send_counts[owners[0]] = partial.num_rows
        else:
            splitters = [boundaries[int(ob[g + 1]) - 1] for g in owners[:-1]]
            cuts = agg.group_cut_points(rt, partial, key, splitters) if splitters else []
            edges = [0] + list(cuts) + [partial.num_rows]
            for d, g in enumerate(owners):
                send_counts[g] = max(0, edges[d + 1] - edges[d])
        recv = coll.exchange_table(rt, partial, send_counts, group)
        # ---- reduce: merge the partials of this rank's key range (K9), finalize ----
        if my_blocks == 0:
            return DeviceDataset(None, [])
        out = agg.aggregate_table(rt, recv, key, aggs, finalize=True) if recv.num_rows else None
        local_bounds = list(boundaries[int(ob[rank]):int(ob[rank + 1]) - 1])
        return DeviceDataset(out, cut_ranges(out, local_bounds, my_blocks))

    def unique_blocks(self, blocks: Sequence[pa.Table], column: str) -> list:
        """``Dataset.unique(column)`` (python/ray/data/dataset.py:2182-2221) over host blocks on
        this rank: the distinct values of the column, ascending, null last."""
        from .. import aggregate as agg

        nonempty = [b.select([column]) for b in blocks if b.num_rows > 0 and b.num_columns > 0]
        if not nonempty:
            return []
        ds = self.stage(nonempty)
        col = agg.unique_values(self.rt, ds.table, column)
        t = dev.export_table(self.rt, dev.DeviceTable([column], [col], col.n, pa.schema([(column, col.type)])))
        return t[column].to_pylist()

    def aggregate_blocks(self, blocks: Sequence[pa.Table], key: Optional[str], aggs: Sequence,
                         boundaries: Optional[list] = None, seed: int = 0, group=None) -> List[pa.Table]:
        """Host blocks in, aggregated host blocks out (this rank's run when distributed).
        Unused columns are pruned before staging (aggregate_task_spec.py:76-109)."""
        from .. import aggregate as agg

        agg.check_aggs(aggs)
        for b in blocks:  # planner/aggregate.py:54-55: aggregates are validated against the schema
            if b.num_columns > 0:
                for a in aggs:
                    a._validate(b.schema)
                if key is not None:
                    SortKey(key).validate_schema(b.schema)
                break
        pruned = [agg.prune_columns(b, agg.key_list(key), aggs) for b in blocks]
        out = self.export(self.aggregate(self.stage(pruned), key, aggs, boundaries, seed, group))
        return [agg.null_typed_like_reference(t) for t in out]

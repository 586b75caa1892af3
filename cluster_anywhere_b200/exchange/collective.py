"""The exchange step between GPUs: one all-to-all-v per column buffer.

In the reference a map task's R output partitions become R objects in the plasma
store and every reduce task pulls its column of the M x R matrix
(python/ray/data/_internal/planner/exchange/pull_based_shuffle_task_scheduler.py:
84-120; transport = src/ray/object_manager, 5 MiB gRPC chunks across nodes).
Here every rank holds its rows already ordered by destination rank, and the
buffers of each column cross NVLink / NVSwitch in a single
``torch.distributed.all_to_all_single`` (NCCL grouped send/recv) with per-peer
split sizes.  Received rows are laid out in (source rank, position at the source)
order -- the (mapper index, in-block position) order the reference's reduce sees.

``torch.distributed`` is plumbing here (process group, collectives); every
conversion of the data itself is a kernel behind the C-ABI.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import pyarrow as pa
import torch
import torch.distributed as dist

from .. import device as dev
from ..device import DeviceColumn, DeviceTable, _vp


# diagnostics: set to a dict to accumulate synchronised wall-clock ms per sub-phase of
# exchange_table (serialises the pipeline; bench.py --trace)
TRACE = None
_t_last = [None]


def _trace(rt, name):
    if TRACE is None:
        return
    import time

    rt.synchronize()
    now = time.perf_counter()
    if name is not None and _t_last[0] is not None:
        TRACE[name] = TRACE.get(name, 0.0) + (now - _t_last[0]) * 1e3
    _t_last[0] = now


def _unwrap(group):
    """-> (comm or None, torch process group).  ``group`` may be a torch process group
    (None = the default one) or one of exchange/comm.py's communicators."""
    from .comm import ThreadComm, TorchComm

    if isinstance(group, ThreadComm):
        return group, None
    if isinstance(group, TorchComm):
        return None, group.group
    return None, group


def world(group=None):
    comm, pg = _unwrap(group)
    if comm is not None:
        return comm.rank, comm.world
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(pg), dist.get_world_size(pg)
    return 0, 1


def all_gather_object(obj, group=None) -> list:
    rank, size = world(group)
    if size == 1:
        return [obj]
    comm, pg = _unwrap(group)
    if comm is not None:
        return comm.all_gather_object(obj)
    out = [None] * size
    dist.all_gather_object(out, obj, group=pg)
    return out


def gather_dataset_info(rt: dev.Runtime, num_blocks: int, num_rows: int, schema, group=None) -> list:
    """[(num_blocks, num_rows, schema)] of every rank.  One small tensor all-gather (block
    count, row count, a process-independent fingerprint of the schema) instead of a
    pickled-object collective (three collectives and two blocking copies: ~10 ms per call
    at 8 ranks, measured); the schemas themselves travel only when a rank has none of its
    own or the fingerprints disagree."""
    rank, size = world(group)
    if size == 1:
        return [(num_blocks, num_rows, schema)]
    import hashlib

    fp = 0
    if schema is not None:
        fp = int.from_bytes(hashlib.blake2b(schema.serialize().to_pybytes(), digest_size=7).digest(), "little") | 1
    mine = torch.tensor([int(num_blocks), int(num_rows), fp], dtype=torch.int64, device=rt.device)
    comm, pg = _unwrap(group)
    if comm is not None:
        rows = comm.all_gather(mine).cpu().numpy()
    else:
        parts = [torch.empty_like(mine) for _ in range(size)]
        dist.all_gather(parts, mine, group=pg)
        rows = torch.stack(parts).cpu().numpy()
    fps = set(int(x) for x in rows[:, 2])
    if len(fps) == 1 and fp != 0:
        return [(int(r[0]), int(r[1]), schema) for r in rows]
    schemas = all_gather_object(schema, group)  # identical decision on every rank
    return [(int(r[0]), int(r[1]), sc) for r, sc in zip(rows, schemas)]


def empty_device_table(rt: dev.Runtime, schema: pa.Schema) -> DeviceTable:
    """Zero-row device table of ``schema`` (a rank may hold no rows but still has to
    take part in the collectives)."""
    cols = []
    for f in schema:
        kind, width = dev._classify(f.type)
        offsets = rt.zeros(1, torch.int64) if kind == "var" else None
        cols.append(DeviceColumn(f.type, 0, kind, width, rt.empty(0), offsets, None, 0))
    return DeviceTable(list(schema.names), cols, 0, schema)


def _a2a(rt: dev.Runtime, send: torch.Tensor, in_splits: Sequence[int], out_splits: Sequence[int],
         dtype=torch.uint8, group=None) -> torch.Tensor:
    recv = rt.empty(int(sum(out_splits)), dtype)
    dist.all_to_all_single(recv, send.contiguous(), [int(x) for x in out_splits],
                           [int(x) for x in in_splits], group=group)
    return recv


def exchange_table(rt: dev.Runtime, table: DeviceTable, send_counts: Sequence[int],
                   group=None) -> DeviceTable:
    """Rows of ``table`` are ordered by destination rank; ``send_counts[d]`` of them go
    to rank ``d``.  Returns the rows this rank receives, in (source rank, position)
    order.  Every rank must call this with the same schema."""
    rank, G = world(group)
    group = _unwrap(group)[1]
    assert len(send_counts) == G
    n_send = table.num_rows
    assert sum(send_counts) == n_send
    cum = np.concatenate([[0], np.cumsum(np.asarray(send_counts, dtype=np.int64))])
    ncols = len(table.columns)
    _trace(rt, None)

    # ---- metadata all-to-all: rows, per-column bytes, validity / uniform flags ----
    # row d of `meta` goes to rank d: [rows, (bytes, has_validity, uniform) per column]
    meta = np.zeros((G, 1 + 3 * ncols), dtype=np.int64)
    meta[:, 0] = send_counts
    cum_d = None
    for j, c in enumerate(table.columns):
        meta[:, 2 + 3 * j] = 1 if c.validity is not None else 0
        if c.kind == "var":
            if c.uniform is not None:
                meta[:, 1 + 3 * j] = np.asarray(send_counts, dtype=np.int64) * c.uniform
                meta[:, 3 + 3 * j] = c.uniform
            else:
                if cum_d is None:
                    cum_d = torch.from_numpy(cum).to(rt.device)
                ends = c.offsets[cum_d].tolist()  # sync: G+1 scalars
                meta[:, 1 + 3 * j] = np.diff(np.asarray(ends, dtype=np.int64))
                meta[:, 3 + 3 * j] = -1
    meta_d = torch.from_numpy(meta.reshape(-1)).to(rt.device)
    got_d = torch.empty_like(meta_d)
    dist.all_to_all_single(got_d, meta_d, group=group)
    got = got_d.cpu().numpy().reshape(G, 1 + 3 * ncols)
    recv_counts = got[:, 0].astype(np.int64)
    n_recv = int(recv_counts.sum())
    rcum = np.concatenate([[0], np.cumsum(recv_counts)])
    _trace(rt, "a2a.meta")

    # a column is exchanged as uniform-width only if every rank sees the same width
    flags = all_gather_object(
        [(int(c.uniform) if (c.kind == "var" and c.uniform is not None) else -1,
          c.validity is not None) for c in table.columns], group)
    _trace(rt, "a2a.flags_gather_object")

    out_cols: List[DeviceColumn] = []
    for j, c in enumerate(table.columns):
        any_validity = any(f[j][1] for f in flags)
        uni = {f[j][0] for f in flags}
        uniform = uni.pop() if (c.kind == "var" and len(uni) == 1 and -1 not in uni) else None

        validity, null_count = None, 0
        if any_validity:
            vb = rt.empty(n_send)
            if c.validity is not None:
                rt.call("xchg_bitmap_unpack", _vp(c.validity), 0, n_send, _vp(vb), rt.stream)
            else:
                vb.fill_(1)
            rb = _a2a(rt, vb, send_counts, recv_counts, group=group)
            validity = rt.zeros(((n_recv + 31) // 32) * 4)
            cnt = rt.zeros(1, torch.int64)
            rt.call("xchg_bitmap_pack", _vp(rb), n_recv, _vp(validity), _vp(cnt), rt.stream)
            null_count = cnt

        if c.kind == "fixed":
            w = c.width
            data = _a2a(rt, c.data[: n_send * w], [int(x) * w for x in send_counts],
                        [int(x) * w for x in recv_counts], group=group)
            out_cols.append(DeviceColumn(c.type, n_recv, "fixed", w, data, None, validity, null_count))
        elif c.kind == "bits":
            vb = rt.empty(n_send)
            rt.call("xchg_bitmap_unpack", _vp(c.data), 0, n_send, _vp(vb), rt.stream)
            rb = _a2a(rt, vb, send_counts, recv_counts, group=group)
            data = rt.zeros(((n_recv + 31) // 32) * 4)
            rt.call("xchg_bitmap_pack", _vp(rb), n_recv, _vp(data), None, rt.stream)
            out_cols.append(DeviceColumn(c.type, n_recv, "bits", 0, data, None, validity, null_count))
        elif uniform is not None:
            w = uniform
            data = _a2a(rt, c.data[: n_send * w], [int(x) * w for x in send_counts],
                        [int(x) * w for x in recv_counts], group=group)
            out_cols.append(DeviceColumn(c.type, n_recv, "var", c.width, data, None, validity,
                                         null_count, uniform=w))
        else:
            offsets = c.offsets
            if offsets is None:  # uniform here, ragged elsewhere: materialise the offsets
                offsets = rt.empty(n_send + 1, torch.int64)
                rt.call("xchg_offsets_uniform64", n_send + 1, c.uniform, 0, _vp(offsets), rt.stream)
                send_bytes = [int(x) * c.uniform for x in send_counts]
            else:
                send_bytes = [int(x) for x in meta[:, 1 + 3 * j]]
            recv_bytes = [int(x) for x in got[:, 1 + 3 * j]]
            total_send = int(sum(send_bytes))
            data = _a2a(rt, c.data[:total_send], send_bytes, recv_bytes, group=group)
            roff = _a2a(rt, offsets[:n_send], send_counts, recv_counts, torch.int64, group=group)
            out_off = rt.empty(n_recv + 1, torch.int64)
            base = 0
            for s in range(G):
                k = int(recv_counts[s])
                if k:
                    rt.call("xchg_offsets_rebase", _vp(roff[int(rcum[s]):]), k, base,
                            _vp(out_off[int(rcum[s]):]), rt.stream)
                base += recv_bytes[s]
            rt.call("xchg_offsets_uniform64", 1, 0, base, _vp(out_off[n_recv:]), rt.stream)
            out_cols.append(DeviceColumn(c.type, n_recv, "var", c.width, data, out_off, validity,
                                         null_count))
        _trace(rt, f"a2a.col{j}.{c.kind}")
    return DeviceTable(list(table.names), out_cols, n_recv, table.schema)

"""Fused take + exchange over NVLink peer memory.

The NCCL path (``collective.exchange_table``) first gathers every column into a
send buffer ordered by destination (K5: read rows at random, write them densely),
then ``ncclSend/ncclRecv`` read that buffer again and write the peer's receive
buffer.  Here the gather kernels write their dense output stream DIRECTLY into the
destination rank's receive buffer: every rank maps one symmetric arena of all its
peers (CUDA VMM handles exchanged once through ``torch.distributed._symmetric_memory``;
plumbing), and K5 is launched once per (column, destination) with the peer's pointer
as its output.  The send buffer, its write and its re-read disappear, and the row
gather runs at link speed with its random HBM reads overlapped tile by tile with the
NVLink stores (a warp's stores are contiguous 128-byte lines, which is what NVLink
moves efficiently).

In the reference this step is M x R plasma objects pulled by the reducers
(python/ray/data/_internal/planner/exchange/pull_based_shuffle_task_scheduler.py:
84-120).

Synchronisation: all arithmetic on where things go is derived from ONE all-gathered
count matrix, identically on every rank; a stream-ordered collective before the
stores orders them after every peer's previous use of its arena, and a second one
after them (or, in the chunked exchange, one arrival flag per chunk) makes the received
rows visible.

Destinations are VIRTUAL: destination v belongs to rank ``vrank[v]`` and owns a row
region of that rank's receive columns; a rank's regions follow one another in the order
of v, each holding the sources' rows in source-rank order.  With one destination per
rank this is the plain all-to-all; the chunked sort gives every rank several (its key
range cut into chunks) so that a chunk can be consumed while later ones are in flight.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import device as dev
from ..device import DeviceColumn, DeviceTable, _vp
from .comm import PeerArena, as_comm

_ALIGN = 256
# start of every arena: arrival flags of the chunked exchange + its wait-timeout word
# (exchange/chunked_sort.py); reserved by every plan so that no column region ever aliases them
HEADER_BYTES = 8192


def _up(x: int) -> int:
    return (int(x) + _ALIGN - 1) // _ALIGN * _ALIGN


def enabled(rt: dev.Runtime, comm=None) -> bool:
    """Peer-memory exchange available: a CUDA device (or a virtual world, whose "peers" live
    in this process) and not switched off by XCHG_P2P=0."""
    if os.environ.get("XCHG_P2P", "1") == "0":
        return False
    return rt.is_cuda or bool(getattr(comm, "virtual", False))


def _ptr(base: int, off: int = 0):
    return ctypes.c_void_p(int(base) + int(off))


class ExchangePlan:
    """Everything every rank derives from the all-gathered count matrix: who sends how
    many rows / bytes to which destination, where each column's region sits in every
    rank's arena."""

    def __init__(self):
        self.rank = 0
        self.G = 1
        self.V = 1
        self.vrank: List[int] = []          # destination -> rank that owns it
        self.table: Optional[DeviceTable] = None
        self.perm: Optional[torch.Tensor] = None
        self.send_counts: List[int] = []    # per destination
        self.cuts = None                    # prefix of send_counts (positions in perm)
        self.M = None                       # [source, destination, 1 + 3 * ncols]
        self.recv_rows = None               # per rank
        self.vrows = None                   # rows of every destination's region
        self.vstart = None                  # first row of the region inside its rank's columns
        self.prefix_rows = None             # [source, destination]: row where the source's run starts
        self.prefix_bytes = {}              # ragged column j -> [source, destination] byte position
        self.col_plan: list = []
        self.offsets_of: list = []
        self.out_offs: dict = {}
        self.arena: Optional[PeerArena] = None
        self.comm = None
        self.header = 0


def local_meta(rt: dev.Runtime, table: DeviceTable, perm, send_counts, out_offs: dict) -> np.ndarray:
    """[destination, 1 + 3 * ncols]: rows | per column: bytes, has_validity, uniform width."""
    V = len(send_counts)
    n = table.num_rows
    ncols = len(table.columns)
    cuts = np.concatenate([[0], np.cumsum(np.asarray(send_counts, dtype=np.int64))])
    meta = np.zeros((V, 1 + 3 * ncols), dtype=np.int64)
    meta[:, 0] = send_counts
    cum_d = None
    for j, c in enumerate(table.columns):
        meta[:, 2 + 3 * j] = 1 if c.validity is not None else 0
        meta[:, 3 + 3 * j] = -1
        if c.kind == "var":
            if c.uniform is not None:
                meta[:, 1 + 3 * j] = np.asarray(send_counts, dtype=np.int64) * c.uniform
                meta[:, 3 + 3 * j] = c.uniform
            elif n:
                if cum_d is None:
                    cum_d = torch.from_numpy(cuts).to(rt.device)
                ends = out_offs[j][cum_d].tolist()  # sync: V+1 scalars
                meta[:, 1 + 3 * j] = np.diff(np.asarray(ends, dtype=np.int64))
    return meta


def plan_exchange(rt: dev.Runtime, holder, table: DeviceTable, perm: Optional[torch.Tensor],
                  send_counts: Sequence[int], group=None, vrank: Optional[Sequence[int]] = None,
                  M: Optional[np.ndarray] = None) -> ExchangePlan:
    """Counts all-gather (a stream-ordered collective: it also orders the stores that
    follow after every peer's previous use of its arena), arena layout, arena
    (re)allocation.  ``M`` = the all-gathered matrix when the caller has it already."""
    comm = as_comm(group)
    rank, G = comm.rank, comm.world
    vrank = list(range(G)) if vrank is None else [int(x) for x in vrank]
    V = len(vrank)
    assert len(send_counts) == V
    n = table.num_rows
    ncols = len(table.columns)
    cuts = np.concatenate([[0], np.cumsum(np.asarray(send_counts, dtype=np.int64))])
    assert int(cuts[-1]) == n
    pl = ExchangePlan()
    pl.rank, pl.G, pl.V, pl.vrank = rank, G, V, vrank
    pl.table, pl.perm, pl.send_counts, pl.cuts, pl.comm = table, perm, list(send_counts), cuts, comm
    pl.header = HEADER_BYTES

    # ---- ragged columns: output offsets of the whole take (needed for byte counts) ----
    for j, c in enumerate(table.columns):
        if c.kind == "var" and c.uniform is None and n:
            oo = rt.empty(n + 1, torch.int64)
            ws_bytes = ctypes.c_uint64()
            dev._cabi.check(dev._cabi.load().xchg_gather_offsets_workspace_bytes(n, ctypes.byref(ws_bytes)))
            ws = rt.empty(ws_bytes.value)
            rt.call("xchg_gather_offsets", _vp(c.offsets), _vp(perm), n, _vp(oo), _vp(ws), ws_bytes.value,
                    rt.stream, launches=3)
            pl.out_offs[j] = oo

    # ---- one all-gathered matrix: [rows | per column: bytes, has_validity, uniform] per destination ----
    if M is None:
        meta = local_meta(rt, table, perm, send_counts, pl.out_offs)
        mine = torch.from_numpy(meta.reshape(-1)).to(rt.device)
        M = comm.all_gather(mine).cpu().numpy().reshape(G, V, 1 + 3 * ncols)
    rows_sv = M[:, :, 0]                                   # [source, destination]
    pl.M = M
    pl.vrows = rows_sv.sum(axis=0)
    vr = np.asarray(vrank)
    pl.recv_rows = np.array([int(pl.vrows[vr == d].sum()) for d in range(G)], dtype=np.int64)
    # a rank's regions follow one another in destination order
    pl.vstart = np.zeros(V, dtype=np.int64)
    for d in range(G):
        acc = 0
        for v in np.nonzero(vr == d)[0]:
            pl.vstart[v] = acc
            acc += int(pl.vrows[v])
    pl.prefix_rows = pl.vstart[None, :] + np.cumsum(rows_sv, axis=0) - rows_sv

    # ---- arena layout of every rank (identical arithmetic everywhere) ----
    pl.offsets_of = [dict() for _ in range(G)]  # rank -> {(col, region): byte offset}
    totals = [pl.header] * G
    for j, c in enumerate(table.columns):
        any_valid = bool(M[:, :, 2 + 3 * j].max() > 0) if V else False
        uni = set(int(x) for x in M[:, :, 3 + 3 * j].reshape(-1))
        uniform = uni.pop() if (c.kind == "var" and len(uni) == 1 and -1 not in uni) else None
        pl.col_plan.append((any_valid, uniform))
        ragged = c.kind == "var" and uniform is None
        if ragged:
            bytes_sv = M[:, :, 1 + 3 * j]
            vbytes = bytes_sv.sum(axis=0)
            bstart = np.zeros(V, dtype=np.int64)
            for d in range(G):
                acc = 0
                for v in np.nonzero(vr == d)[0]:
                    bstart[v] = acc
                    acc += int(vbytes[v])
            pl.prefix_bytes[j] = bstart[None, :] + np.cumsum(bytes_sv, axis=0) - bytes_sv
        for d in range(G):
            r = int(pl.recv_rows[d])
            if any_valid:
                pl.offsets_of[d][(j, "valid")] = totals[d]
                totals[d] += _up(r)
            if c.kind == "fixed":
                nb = r * c.width
            elif c.kind == "bits":
                nb = r
            elif uniform is not None:
                nb = r * uniform
            else:
                nb = int(M[:, vr == d, 1 + 3 * j].sum())
                pl.offsets_of[d][(j, "offsets")] = totals[d]
                totals[d] += _up((r + 1) * 8)
            pl.offsets_of[d][(j, "data")] = totals[d]
            totals[d] += _up(nb + 64)
    need = max(totals) + _ALIGN
    arena = getattr(holder, "arena", None)
    if arena is None or arena.nbytes < need:
        if arena is not None:
            holder.arena = None
            del arena
        arena = comm.new_arena(rt, int(need * 1.25) + (1 << 20))
        arena.buf[: pl.header].zero_()  # arrival flags start at zero
        rt.synchronize()
        comm.barrier()
        holder.arena = arena
        holder.arena_epoch = 0
    pl.arena = arena
    return pl


def store_columns(rt: dev.Runtime, pl: ExchangePlan, cols: Sequence[int],
                  vsel: Optional[Sequence[int]] = None) -> None:
    """K5 once per (column, destination), output = the owning rank's arena (current stream).
    ``vsel`` = the destinations to store now (default: all)."""
    table, perm, arena, M, rank, G = pl.table, pl.perm, pl.arena, pl.M, pl.rank, pl.G
    perm_ptr = perm.data_ptr() if perm is not None else 0
    vsel = list(range(pl.V)) if vsel is None else list(vsel)
    # destinations in rotated rank order (rank+1, rank+2, ..., rank): at any moment every
    # rank stores into a different peer, so each GPU's NVLink ingress serves one
    # sender at a time instead of all G-1 at once (in plain 0..G-1 order every rank
    # hits destination 0 together, then 1, ...: an 8-GPU step took 95 ms)
    order = sorted(vsel, key=lambda v: ((pl.vrank[v] - rank - 1) % G, v))
    for j in cols:
        c = table.columns[j]
        any_valid, uniform = pl.col_plan[j]
        multi = (c.kind == "var" and uniform is not None and uniform > 0
                 and os.environ.get("XCHG_MULTI_STORE", "1") == "1")
        if multi:
            # ONE launch storing to every destination at once (tiles of the destinations
            # interleaved): each peer's ingress is shared evenly by all senders throughout
            dests = [v for v in order if int(pl.send_counts[v]) > 0]
            for g0 in range(0, len(dests), 16):
                part = dests[g0:g0 + 16]
                row0 = (ctypes.c_uint64 * len(part))(*[int(pl.cuts[v]) for v in part])
                rows = (ctypes.c_uint64 * len(part))(*[int(pl.send_counts[v]) for v in part])
                outs = (ctypes.c_void_p * len(part))(*[
                    int(arena.ptrs[pl.vrank[v]]) + pl.offsets_of[pl.vrank[v]][(j, "data")]
                    + int(pl.prefix_rows[rank, v]) * uniform for v in part])
                try:
                    rt.call("xchg_gather_rows_multi", _vp(c.data), uniform, _ptr(perm_ptr), len(part), row0, rows,
                            outs, rt.stream)
                except dev.XchgError:
                    multi = False  # rows wider than the staging tile: one launch per destination
                    break
        for v in order:
            k = int(pl.send_counts[v])
            if k == 0:
                continue
            d = pl.vrank[v]
            lo = int(pl.cuts[v])
            idx = _ptr(perm_ptr, lo * 4)
            pre = int(pl.prefix_rows[rank, v])
            base = arena.ptrs[d]
            off = pl.offsets_of[d]
            if any_valid:
                dst = _ptr(base, off[(j, "valid")] + pre)
                if c.validity is not None:
                    rt.call("xchg_gather_bitmap_bytes", _vp(c.validity), 0, idx, k, dst, rt.stream)
                else:
                    ones = getattr(pl, "_ones", None)
                    if ones is None or ones.numel() < k:
                        ones = torch.ones(max(k, int(max(pl.send_counts))), dtype=torch.uint8, device=rt.device)
                        pl._ones = ones
                    src = (ctypes.c_void_p * 1)(ones.data_ptr())
                    dsts = (ctypes.c_void_p * 1)(dst.value)
                    nb = (ctypes.c_uint64 * 1)(k)
                    rt.call("xchg_copy_segments", 1, src, dsts, nb, rt.stream)
            data_off = off[(j, "data")]
            if c.kind == "fixed":
                rt.call("xchg_gather_fixed", _vp(c.data), c.width, idx, k, _ptr(base, data_off + pre * c.width),
                        rt.stream)
            elif c.kind == "bits":
                rt.call("xchg_gather_bitmap_bytes", _vp(c.data), 0, idx, k, _ptr(base, data_off + pre), rt.stream)
            elif uniform is not None:
                if not multi:
                    rt.call("xchg_gather_rows", _vp(c.data), uniform, idx, k, _ptr(base, data_off + pre * uniform),
                            rt.stream)
            else:
                pre_bytes = int(pl.prefix_bytes[j][rank, v])
                if c.uniform is not None:  # uniform here, ragged elsewhere
                    w = c.uniform
                    rt.call("xchg_gather_rows", _vp(c.data), w, idx, k, _ptr(base, data_off + pre_bytes), rt.stream)
                    rt.call("xchg_offsets_uniform64", k, w, pre_bytes,
                            _ptr(base, off[(j, "offsets")] + pre * 8), rt.stream)
                else:
                    oo = pl.out_offs[j]
                    first = int(M[rank, :v, 1 + 3 * j].sum())  # = oo[lo]: bytes of earlier destinations
                    # row i of the segment lands at  data + pre_bytes + (oo[lo+i] - oo[lo])
                    rt.call("xchg_gather_bytes", _vp(c.data), _vp(c.offsets), idx, k, _ptr(oo.data_ptr(), lo * 8),
                            _ptr(base, data_off + pre_bytes - first), rt.stream)
                    rt.call("xchg_offsets_rebase", _ptr(oo.data_ptr(), lo * 8), k, pre_bytes,
                            _ptr(base, off[(j, "offsets")] + pre * 8), rt.stream)


def fence(rt: dev.Runtime, group=None) -> None:
    """After this stream-ordered collective every rank's preceding stores (on the stream
    the collective was issued from) are complete and visible."""
    as_comm(group).fence(rt)


def received_columns(rt: dev.Runtime, pl: ExchangePlan, cols: Sequence[int]) -> List[DeviceColumn]:
    """Views of the local arena for the given columns (valid until the next exchange)."""
    table, arena, M, rank = pl.table, pl.arena, pl.M, pl.rank
    m = int(pl.recv_rows[rank])
    my = pl.offsets_of[rank]
    out_cols: List[DeviceColumn] = []
    for j in cols:
        c = table.columns[j]
        any_valid, uniform = pl.col_plan[j]
        validity, nulls = None, 0
        if any_valid and m:
            vb = arena.buf[my[(j, "valid")]: my[(j, "valid")] + m]
            validity = rt.zeros(((m + 31) // 32) * 4)
            cnt = rt.zeros(1, torch.int64)
            rt.call("xchg_bitmap_pack", _vp(vb), m, _vp(validity), _vp(cnt), rt.stream)
            nulls = cnt
        o = my[(j, "data")]
        if c.kind == "fixed":
            out_cols.append(DeviceColumn(c.type, m, "fixed", c.width, arena.buf[o: o + m * c.width], None,
                                         validity, nulls))
        elif c.kind == "bits":
            data = rt.zeros(((m + 31) // 32) * 4)
            if m:
                rt.call("xchg_bitmap_pack", _vp(arena.buf[o: o + m]), m, _vp(data), None, rt.stream)
            out_cols.append(DeviceColumn(c.type, m, "bits", 0, data, None, validity, nulls))
        elif uniform is not None:
            out_cols.append(DeviceColumn(c.type, m, "var", c.width, arena.buf[o: o + m * uniform], None,
                                         validity, nulls, uniform=uniform))
        else:
            total = int(M[:, np.asarray(pl.vrank) == rank, 1 + 3 * j].sum())
            oo_off = my[(j, "offsets")]
            offs = arena.buf[oo_off: oo_off + (m + 1) * 8].view(torch.int64)
            rt.call("xchg_offsets_uniform64", 1, 0, total, _ptr(offs.data_ptr(), m * 8), rt.stream)
            out_cols.append(DeviceColumn(c.type, m, "var", c.width, arena.buf[o: o + total], offs, validity, nulls))
    return out_cols


def gather_exchange(rt: dev.Runtime, holder, table: DeviceTable, perm: Optional[torch.Tensor],
                    send_counts: Sequence[int], group=None) -> DeviceTable:
    """``take(table, perm)`` ordered by destination rank, written straight into the
    destinations' arenas.  ``perm`` (uint32 row ids, may be None when there are no rows)
    lists this rank's rows by destination; ``send_counts[d]`` of them go to rank ``d``.
    ``holder.arena`` caches the symmetric arena between calls.  Returns the received
    rows in (source rank, position at the source) order; the table aliases the arena and
    is valid until the next exchange."""
    pl = plan_exchange(rt, holder, table, perm, send_counts, group)
    every = list(range(len(table.columns)))
    store_columns(rt, pl, every)
    fence(rt, group)
    cols = received_columns(rt, pl, every)
    return DeviceTable(list(table.names), cols, int(pl.recv_rows[pl.rank]), table.schema)

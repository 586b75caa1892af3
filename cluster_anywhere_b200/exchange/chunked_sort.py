"""Multi-GPU ``Dataset.sort`` fast path: device-side sampling, one stable partition pass,
a chunked peer-store exchange with arrival flags, and a receive side that sorts and
gathers chunk c while chunks c+1.. are still crossing NVLink.

The reference pipelines its shuffle the same way one level up: rounds of map tasks
overlap with rounds of merge tasks that consume the previous round's outputs
(python/ray/data/_internal/planner/exchange/push_based_shuffle_task_scheduler.py:548-561,
merge with ``partial_reduce=True`` :696-730), and the final reduce of partition j starts
when partition j's merge outputs exist (:330-397).  Here a "round" is a CHUNK: rank g's
key range (its run of output blocks) is cut at block boundaries into up to
``XCHG_CHUNKS`` sub-ranges; every rank partitions its rows by (destination rank, chunk)
in one stable pass, stores the payload chunk after chunk into the owners' arenas and
raises one 4-byte arrival flag per (owner, chunk); the owner's final gather of chunk c
waits for the G flags of chunk c, never for the whole exchange.  The radix keys (8 % of
the bytes) go first, in one piece: the owner sorts them once while the payload is still
crossing NVLink.

Per rank and step (device-resident input):

  main stream   K1 sample -> all-gather -> [D2H of 8 header words per rank]
                boundary selection (one CTA) -> destination histogram -> all-gather of
                the counts -> stable partition (radix key, row id) by destination
                [D2H of the count matrix] -> key runs copied to the owners, flag per chunk
  side stream   for every chunk c: payload rows gathered straight into the owners'
                arenas (K5 storing over NVLink, a CTA or two per SM), flag per chunk
  receive stream  wait(keys) -> ONE stable sort of all received radix keys -> block cuts
                (K4) -> key column from the sorted radix keys; then for every own chunk c:
                wait(payload, c) -> final payload gather (K5) of the chunk's slice of the
                sorted order

Eligible: ONE fixed-width key column without nulls (any of the ten key types;
descending floats excepted), payload columns fixed-width or uniform-width without
nulls, boundaries sampled or given as plain numbers.  Everything else takes the general
path (``DeviceExchange.sort_distributed``).  Both produce the same rows: the global
stable order with ties in (rank, row) = (block, position) order.
"""
from __future__ import annotations

import ctypes
import hashlib
import os
from typing import List, Optional

import numpy as np
import pyarrow as pa
import torch

from .. import _cabi
from .. import device as dev
from .. import keys as keymod
from ..device import DeviceColumn, DeviceTable, _vp
from ..sort_key import SortKey
from . import peer

HDR = 8            # int64 words: blocks, rows, schema fingerprint, samples, eligible, layout signature
CAP = 4096         # sample slots per rank in the all-gathered message
# The reference draws int(R * 10 / M) = 10 rows per block when R == M (sort_task_spec.py:167);
# with 10 the gap between two boundaries has a relative spread of 1 / sqrt(10): blocks up to
# ~2.5x the mean, ranks +-10 %.  Sampling is a kernel here and costs nothing, so every block
# gives up to 100 samples (cut points are unpinned by the reference; XCHG_SAMPLES_PER_BLOCK=10
# restores its count).
SAMPLES_PER_BLOCK = int(os.environ.get("XCHG_SAMPLES_PER_BLOCK", "100"))
CMAX = 16          # chunks per rank the flag area has room for
GMAX = 32
FLAG_AREA = 4096   # 2 kinds x CMAX x GMAX uint32
HEADER_BYTES = peer.HEADER_BYTES
ERR_OFF = FLAG_AREA
KIND_KEYS, KIND_PAYLOAD = 0, 1
WAIT_TIMEOUT_NS = int(float(os.environ.get("XCHG_WAIT_TIMEOUT_S", "20")) * 1e9)
# timing experiments only (the output is then wrong): 1 = no payload stores, 2 = no receive side
_DEBUG_SKIP = int(os.environ.get("XCHG_DEBUG_SKIP", "0"))


def _ptr(base: int, off: int = 0):
    return ctypes.c_void_p(int(base) + int(off))


def _layout_signature(table: DeviceTable) -> int:
    h = hashlib.blake2b(digest_size=7)
    for name, c in zip(table.names, table.columns):
        h.update(repr((name, str(c.type), c.kind, c.width, c.uniform, c.validity is not None)).encode())
    return int.from_bytes(h.digest(), "little") | 1


def _schema_fp(schema: pa.Schema) -> int:
    return int.from_bytes(hashlib.blake2b(schema.serialize().to_pybytes(), digest_size=7).digest(), "little") | 1


def _eligible(ds, sort_key: SortKey, boundaries) -> bool:
    cols = sort_key.get_columns()
    if len(cols) != 1 or ds.table is None:
        return False
    t = ds.table
    if cols[0] not in t.names:
        return False
    kc = t.column(cols[0])
    if kc.kind != "fixed" or kc.validity is not None:
        return False
    try:
        kt = dev.key_type_of(kc.type)
    except NotImplementedError:
        return False
    if kt in (_cabi.F32, _cabi.F64) and sort_key.get_descending()[0]:
        return False
    for c in t.columns:
        if c.validity is not None:
            return False
        if not (c.kind == "fixed" or (c.kind == "var" and c.uniform is not None and c.uniform > 0)):
            return False
    if ds.num_blocks * 10 > CAP:
        return False
    if boundaries is not None:
        for b in boundaries:
            if not isinstance(b, (tuple, list)) or len(b) != 1:
                return False
            v = b[0]
            if v is None or isinstance(v, (str, bytes)) or (isinstance(v, float) and v != v):
                return False
    return True


def try_sort(ex, ds, sort_key: SortKey, boundaries: Optional[list], seed: int, comm):
    """-> DeviceDataset, or None when some rank's input is not eligible (every rank then
    takes the general path; the decision is collective)."""
    from .device_exchange import DeviceDataset

    rt = ex.rt
    rank, G = comm.rank, comm.world
    if G > GMAX:
        return None
    ok = _eligible(ds, sort_key, boundaries)
    table = ds.table
    desc = bool(sort_key.get_descending()[0]) if sort_key.get_columns() else False
    n = ds.num_rows
    ex._gpu_mark("step_begin")

    # ---- K1 + one all-gather: [header | samples] of every rank ----
    msg = rt.zeros(HDR + CAP, torch.int64)
    nsamp = ctypes.c_uint64(0)
    kt = 0
    if ok:
        keycol = table.column(sort_key.get_columns()[0])
        kt = dev.key_type_of(keycol.type)
        if boundaries is None and n:
            nb = ds.num_blocks
            starts = (ctypes.c_uint64 * nb)(*[lo for lo, hi in ds.block_ranges])
            rows = (ctypes.c_uint64 * nb)(*[max(0, hi - lo) for lo, hi in ds.block_ranges])
            per_block = max(10, min(SAMPLES_PER_BLOCK, CAP // max(1, nb)))
            rt.call("xchg_sample_keys", _vp(keycol.data), kt, int(desc), None, 0, starts, rows, nb,
                    per_block, int(seed) & 0xFFFFFFFFFFFFFFFF, rank << 32,
                    _ptr(msg.data_ptr(), HDR * 8), CAP, ctypes.byref(nsamp), rt.stream)
    hdr = [ds.num_blocks, n, _schema_fp(table.schema) if table is not None else 0, int(nsamp.value),
           int(ok), _layout_signature(table) if table is not None else 0, 0, 0]
    msg[:HDR].copy_(torch.tensor(hdr, dtype=torch.int64), non_blocking=True)
    gathered = comm.all_gather(msg)                       # [G, HDR + CAP]
    hdrs = gathered[:, :HDR].cpu().numpy()                # read-back 1: 8 words per rank
    ex._gpu_mark("g.info_gathered")
    if not (hdrs[:, 4].all() and len(set(hdrs[:, 2].tolist())) == 1 and len(set(hdrs[:, 5].tolist())) == 1):
        return None
    blocks_per_rank = [int(x) for x in hdrs[:, 0]]
    rows_per_rank = [int(x) for x in hdrs[:, 1]]
    total_rows = sum(rows_per_rank)
    r_total = max(1, sum(blocks_per_rank) if boundaries is None else len(boundaries) + 1)
    if boundaries is None and all(b > 0 for b in blocks_per_rank):
        out_per_rank = list(blocks_per_rank)
    else:
        out_per_rank = [r_total // G + (1 if g < r_total % G else 0) for g in range(G)]
    ob = np.concatenate([[0], np.cumsum(out_per_rank)]).astype(int)
    my_blocks = int(out_per_rank[rank])
    if total_rows == 0:
        return DeviceDataset(None, [(0, 0)] * my_blocks)
    sort_key.validate_schema(table.schema)
    names = list(table.names)
    key_j = names.index(sort_key.get_columns()[0])
    keycol = table.columns[key_j]
    int_key = kt not in (_cabi.F32, _cabi.F64)

    # ---- boundaries, resident on the device as radix keys ----
    nbound = r_total - 1
    bflags_d = None
    if boundaries is None:
        bounds_d = rt.empty(max(1, nbound), torch.int64)
        rt.call("xchg_select_boundaries", _vp(gathered), G, HDR + CAP, 3, HDR, r_total, _vp(bounds_d), None,
                rt.stream)
        strict_all = [1 if desc else 0] * nbound
    else:
        lower, _, flags = keymod.normalize_boundaries(boundaries, [kt], [desc])
        bounds_d = rt.empty(max(1, nbound), torch.int64)
        if nbound:
            dev._h2d(bounds_d.view(torch.uint8)[: nbound * 8],
                     np.ascontiguousarray(lower[:, 0], dtype=np.uint64).view(np.uint8))
        fl = np.ascontiguousarray(flags[:, 0], dtype=np.uint8)
        strict_all = [1 if (desc and not (int(f) & keymod.FLAG_EMPTY)) else 0 for f in fl]
        if fl.any():
            bflags_d = rt.empty(max(1, nbound), torch.uint8)
            dev._h2d(bflags_d[:nbound], fl)

    # ---- virtual destinations: (owner rank, chunk of its block run) ----
    cmax = max(1, min(CMAX, int(os.environ.get("XCHG_CHUNKS", "4")), 256 // G))
    owners = [g for g in range(G) if out_per_rank[g] > 0]
    vrank, vblocks, vchunk = [], [], []          # per destination: rank, (first block, end block), chunk index
    chunks_of = {g: [] for g in range(G)}
    for g in owners:
        nbk = out_per_rank[g]
        cg = min(cmax, nbk)
        for c in range(cg):
            b0, b1 = int(ob[g]) + c * nbk // cg, int(ob[g]) + (c + 1) * nbk // cg
            chunks_of[g].append(len(vrank))
            vrank.append(g)
            vblocks.append((b0, b1))
            vchunk.append(c)
    V = len(vrank)
    nchunks = max(len(v) for v in chunks_of.values())
    pick = [vblocks[v][0] - 1 for v in range(1, V)]   # boundary index in front of destination v
    nthr = V - 1
    pick_a = (ctypes.c_uint32 * max(1, nthr))(*pick) if nthr else None
    strict_a = (ctypes.c_uint8 * max(1, nthr))(*[strict_all[p] for p in pick]) if nthr else None

    # ---- destination histogram, counts all-gather, stable partition ----
    counts = rt.zeros(V, torch.int64)
    pk = prow = None
    if n:
        rt.call("xchg_partition_hist", _vp(keycol.data), kt, int(desc), n, _vp(bounds_d), pick_a, strict_a, nthr,
                _vp(counts), rt.stream)
    allc = comm.all_gather(counts)                        # [G, V]
    if n:
        wsb = ctypes.c_uint64()
        _cabi.check(_cabi.load().xchg_partition_workspace_bytes(n, ctypes.byref(wsb)))
        pws = rt.empty(wsb.value)
        pk = rt.empty(n, torch.int64)
        prow = rt.empty(n, torch.int32)
        rt.call("xchg_partition_rows", _vp(keycol.data), kt, int(desc), n, _vp(bounds_d), pick_a, strict_a, nthr,
                _vp(counts), _vp(pk), _vp(prow), _vp(pws), wsb.value, rt.stream, launches=2)
    rows_sv = allc.cpu().numpy().astype(np.int64)         # read-back 2: the count matrix
    ex._gpu_mark("g.assigned")
    send_counts = [int(x) for x in rows_sv[rank]]

    # ---- wire table: payload columns (+ the raw key column when it cannot be rebuilt
    # from radix keys) + the radix keys as one more fixed-width column ----
    pay_j = [j for j in range(len(names)) if not (j == key_j and int_key)]
    wire_cols = [table.columns[j] for j in pay_j]
    wire_names = [names[j] for j in pay_j]
    kcol = DeviceColumn(pa.uint64(), n, "fixed", 8, pk.view(torch.uint8) if pk is not None else rt.empty(0))
    wire = DeviceTable(wire_names + ["__xchg_key"], wire_cols + [kcol], n,
                       pa.schema([table.schema.field(j) for j in pay_j] + [pa.field("__xchg_key", pa.uint64())]))
    ncw = len(wire.columns)
    M = np.zeros((G, V, 1 + 3 * ncw), dtype=np.int64)
    M[:, :, 0] = rows_sv
    for j, c in enumerate(wire.columns):
        M[:, :, 3 + 3 * j] = c.uniform if c.kind == "var" else -1
        if c.kind == "var":
            M[:, :, 1 + 3 * j] = rows_sv * c.uniform
    pl = peer.plan_exchange(rt, ex, wire, prow, send_counts, comm, vrank=vrank, M=M)
    arena = pl.arena
    ex.arena_epoch = getattr(ex, "arena_epoch", 0) + 1
    epoch = ex.arena_epoch
    kj = ncw - 1
    main = torch.cuda.current_stream(rt.device) if rt.is_cuda else None

    def flag_ptrs(kind, c):
        """this rank's flag of (kind, chunk c) in every owner that has a chunk c"""
        tg = [g for g in owners if c < len(chunks_of[g])]
        return (ctypes.c_void_p * len(tg))(*[arena.ptrs[g] + ((kind * CMAX + c) * GMAX + rank) * 4 for g in tg]), len(tg)

    def local_flags(kind, c):
        return _ptr(arena.buf.data_ptr(), ((kind * CMAX + c) * GMAX) * 4)

    err_ptr = _ptr(arena.buf.data_ptr(), ERR_OFF)

    # ---- payload: side stream, chunk after chunk ----
    side = _stream(ex, "_side_stream")
    srt = _stream(ex, "_sort_stream")
    pay_cols = list(range(ncw - 1))

    def push_payload():
        # the stores are link-bound: a CTA or two per SM keep NVLink busy and leave the rest of
        # every SM (shared memory above all) to the receive-side sorts and gathers
        cap = int(os.environ.get("XCHG_SIDE_CTAS_PER_SM", "2"))
        stage_kb = int(os.environ.get("XCHG_SIDE_STAGE_KB", "0"))
        if cap:
            rt.call("xchg_set_gather_ctas_per_sm", cap, launches=0)
        if stage_kb:
            rt.call("xchg_set_gather_stage_kb", stage_kb, launches=0)
        try:
            for c in range(nchunks):
                vs = [chunks_of[g][c] for g in owners if c < len(chunks_of[g])]
                if n and pay_cols and not (_DEBUG_SKIP & 1):
                    peer.store_columns(rt, pl, pay_cols, vsel=vs)
                fp, k = flag_ptrs(KIND_PAYLOAD, c)
                rt.call("xchg_signal_flags", fp, k, epoch, rt.stream)
        finally:
            if cap:
                rt.call("xchg_set_gather_ctas_per_sm", 0, launches=0)
            if stage_kb:
                rt.call("xchg_set_gather_stage_kb", 0, launches=0)

    partitioned = None
    if main is not None:
        partitioned = torch.cuda.Event()
        partitioned.record(main)

    # ---- key runs: contiguous copies into the owners' arenas, ONE flag per owner (queued
    # before the payload so that their CTAs are placed first; the keys are 8 % of the bytes
    # and the receive-side sort wants all of them at once) ----
    srcs, dsts, nbs = [], [], []
    for c in range(nchunks):
        for g in owners:
            if c >= len(chunks_of[g]):
                continue
            v = chunks_of[g][c]
            k = send_counts[v]
            if k:
                srcs.append(pk.data_ptr() + int(pl.cuts[v]) * 8)
                dsts.append(arena.ptrs[g] + pl.offsets_of[g][(kj, "data")] + int(pl.prefix_rows[rank, v]) * 8)
                nbs.append(k * 8)
    if srcs:
        rt.call("xchg_copy_segments", len(srcs), (ctypes.c_void_p * len(srcs))(*srcs),
                (ctypes.c_void_p * len(srcs))(*dsts), (ctypes.c_uint64 * len(srcs))(*nbs), rt.stream,
                launches=(len(srcs) + 15) // 16)
    fp, k = flag_ptrs(KIND_KEYS, 0)
    rt.call("xchg_signal_flags", fp, k, epoch, rt.stream)
    ex._gpu_mark("g.keys_stored")

    # ---- payload: side stream, chunk after chunk ----
    if side is not None:
        with torch.cuda.stream(side):
            side.wait_event(partitioned)
            push_payload()
            ex._gpu_mark("g.side_payload_stored")
    else:  # host double: no streams, everything in program order
        push_payload()

    # ---- receive: chunk by chunk ----
    m = int(pl.recv_rows[rank])
    mine = chunks_of[rank]
    out_ds = None
    cuts_d = None
    lo_b, hi_b = int(ob[rank]), int(ob[rank + 1])
    nb_local = max(0, hi_b - 1 - lo_b)
    if my_blocks and m:
        my_off = pl.offsets_of[rank]
        norm_sorted = rt.empty(m, torch.int64)
        out_cols = {}
        for jj, j in enumerate(pay_j):
            c = wire.columns[jj]
            w = c.width if c.kind == "fixed" else c.uniform
            out_cols[j] = (c, w, rt.empty(m * w))
        out_key = rt.empty(m * keycol.width) if int_key else None
        ws_bytes = dev.sort_workspace_bytes(m)
        ws = rt.empty(ws_bytes)
        if nb_local:
            cuts_d = rt.zeros(nb_local, torch.int64)
            thr_last, has_last = keymod.last_class_threshold(kt, 0)
        perm = rt.empty(m, torch.int32)

        def sort_keys():
            """ONE stable sort of all received radix keys (positions in arena order as the
            values): the chunks are consecutive key ranges, so chunk c's rows end up in the
            sorted positions [vstart_c, vstart_c + rows_c) and its final gather needs nothing
            but that slice of the permutation."""
            rt.call("xchg_wait_flags", local_flags(KIND_KEYS, 0), G, epoch, WAIT_TIMEOUT_NS, err_ptr, rt.stream)
            kin = _ptr(arena.buf.data_ptr(), my_off[(kj, "data")])
            rt.call("xchg_sort_column", kin, _cabi.U64, 0, None, 0, None, m, _vp(perm), _vp(norm_sorted), _vp(ws),
                    ws_bytes, rt.stream, launches=8)
            if nb_local:
                ptrs = (ctypes.c_void_p * 1)(norm_sorted.data_ptr())
                rt.call("xchg_partition_bounds", ptrs, (ctypes.c_uint8 * 1)(int(desc)),
                        (ctypes.c_uint64 * 1)(int(thr_last)), (ctypes.c_uint8 * 1)(int(has_last)), 1, m,
                        _ptr(bounds_d.data_ptr(), lo_b * 8), _ptr(bounds_d.data_ptr(), lo_b * 8),
                        _ptr(bflags_d.data_ptr(), lo_b) if bflags_d is not None else None,
                        nb_local, int(desc), _vp(cuts_d), rt.stream)
            if int_key:
                rt.call("xchg_denormalize_keys", _vp(norm_sorted), kt, int(desc), m, _vp(out_key), rt.stream)
            ex._gpu_mark("g.keys_sorted")

        def gather_chunk(ci):
            v = mine[ci]
            mc, base = int(pl.vrows[v]), int(pl.vstart[v])
            rt.call("xchg_wait_flags", local_flags(KIND_PAYLOAD, ci), G, epoch, WAIT_TIMEOUT_NS, err_ptr, rt.stream)
            if mc:
                idx = _ptr(perm.data_ptr(), base * 4)
                for jj, j in enumerate(pay_j):
                    c, w, out = out_cols[j]
                    src = _ptr(arena.buf.data_ptr(), my_off[(jj, "data")])
                    dst = _ptr(out.data_ptr(), base * w)
                    if c.kind == "fixed":
                        rt.call("xchg_gather_fixed", src, w, idx, mc, dst, rt.stream)
                    else:
                        rt.call("xchg_gather_rows", src, w, idx, mc, dst, rt.stream)
            ex._gpu_mark(f"g.taken_{ci}")

        # ONE receive stream: the sort and the final gathers each want the whole SM (shared
        # memory), so they take turns; the payload stores keep their few resident CTAs per SM
        # throughout.  The keys (8 % of the bytes) are all there long before the payload, so
        # the sort runs under the payload's transfer and the gathers follow the chunks in.
        import contextlib

        if srt is not None:
            srt.wait_stream(main)
        with (torch.cuda.stream(srt) if srt is not None else contextlib.nullcontext()):
            if not (_DEBUG_SKIP & 2):
                sort_keys()
                for ci in range(len(mine)):
                    gather_chunk(ci)
        cols = []
        for j in range(len(names)):
            src = table.columns[j]
            if j == key_j and int_key:
                cols.append(DeviceColumn(src.type, m, "fixed", src.width, out_key))
            else:
                c, w, out = out_cols[j]
                if c.kind == "fixed":
                    cols.append(DeviceColumn(src.type, m, "fixed", src.width, out))
                else:
                    cols.append(DeviceColumn(src.type, m, "var", src.width, out, None, None, 0, uniform=w))
        out_ds = DeviceTable(names, cols, m, table.schema)
    if main is not None:
        main.wait_stream(srt)
        main.wait_stream(side)
    ex._gpu_mark("g.taken")

    # ---- read-back 3 (after every launch): block cuts and the wait-timeout flag ----
    err = int(arena.buf[ERR_OFF: ERR_OFF + 4].view(torch.int32).item())
    if err:
        arena.buf[ERR_OFF: ERR_OFF + 4].zero_()
        raise dev.XchgError(f"multi-GPU exchange: arrival flag {err - 1} did not arrive within "
                            f"{WAIT_TIMEOUT_NS / 1e9:.0f} s (a peer rank failed or stalled)")
    ex.chunked_steps = getattr(ex, "chunked_steps", 0) + 1
    if not my_blocks:
        return DeviceDataset(None, [])
    if out_ds is None:
        return DeviceDataset(None, [(0, 0)] * my_blocks)
    cuts: List[int] = []
    if nb_local:
        cuts = [int(x) for x in cuts_d.cpu().numpy()]
    ranges, last = [], 0
    for c in cuts:
        ranges.append((last, c))
        last = c
    ranges.append((last, m))
    return DeviceDataset(out_ds, ranges)


def _stream(ex, attr: str):
    if not ex.rt.is_cuda:
        return None
    s = getattr(ex, attr, None)
    if s is None:
        # CTAs of a later kernel are placed only when the earlier kernels have none left to
        # place, unless its stream has priority: the short receive-side sorts (which the final
        # gathers wait for) go first, then the payload stores (the long pole whenever NVLink
        # binds), the final gathers take what is left.  XCHG_STREAM_PRIORITIES=0: all equal.
        prio = {"_sort_stream": -2, "_side_stream": -1}.get(attr, 0)
        if os.environ.get("XCHG_STREAM_PRIORITIES", "1") == "0":
            prio = 0
        s = torch.cuda.Stream(device=ex.rt.device, priority=prio)
        setattr(ex, attr, s)
    return s

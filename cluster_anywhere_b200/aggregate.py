"""Aggregations of ``groupby().aggregate()`` and their device implementation.

Host surface: ``Count / Sum / Min / Max / Mean / Std`` with the reference's
constructor arguments, result names and null policy (python/ray/data/aggregate.py:
97-326, _internal/null_aggregate.py:11-25).  ``AbsMax / Quantile / Unique`` are row /
list based in the reference (:330-489) and stay on the reference path: asking this
engine for them raises ``NotImplementedError``.

Device side (``aggregate_table``): the hash group-by of csrc/aggregate.cu (K8), its
partial-merge twin (K9), and the finalisation into the reference's output layout --
key column first (integers -> int64, floats -> double), one column per aggregate,
rows sorted by key with NaN then null last.

Partial blocks (what ``combine`` emits and ``aggregate_combined_blocks`` merges --
the reference's accumulator columns, an internal format): ``[key, __xchg_rows,
__xchg_state:<col> ...]`` where a state is 6 eight-byte words
``cnt, sum, min, max, nan, M2`` stored as ``fixed_size_binary(48)``; the value
column's type travels in the field metadata.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import torch

from . import _cabi
from . import device as dev
from .device import DeviceColumn, DeviceTable, _vp

KIND_COUNT, KIND_SUM, KIND_MIN, KIND_MAX, KIND_MEAN, KIND_STD = range(6)
ROWS_COL = "__xchg_rows"
STATE_PREFIX = "__xchg_state:"
STATE_WORDS = 6
_VT_META = b"xchg_value_type"


class AggregateFn:
    """Base of the built-in aggregations (aggregate.py:20-81).  The engine does not run
    user callbacks on the device: only the built-in kinds below are accepted."""

    kind: int = -1

    def __init__(self, name: str, on: Optional[str] = None, ignore_nulls: bool = True, ddof: int = 1):
        if not isinstance(name, str):
            raise TypeError("`name` must be provided.")
        self.name = name
        self._key_fn = on
        self.on = on
        self.ignore_nulls = bool(ignore_nulls)
        self.ddof = int(ddof)

    def _validate(self, schema) -> None:
        """aggregate.py:83-94: an aggregate on a column checks that the column exists."""
        if self.on is not None:
            from .sort_key import SortKey

            SortKey(self.on).validate_schema(schema)


class Count(AggregateFn):
    """aggregate.py:97-108: number of rows of the group (nulls included)."""

    kind = KIND_COUNT

    def __init__(self):
        super().__init__("count()")


def _named(prefix: str, on, alias_name):
    return alias_name if alias_name else f"{prefix}({str(on)})"


class Sum(AggregateFn):
    kind = KIND_SUM

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        super().__init__(_named("sum", on, alias_name), on, ignore_nulls)


class Min(AggregateFn):
    kind = KIND_MIN

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        super().__init__(_named("min", on, alias_name), on, ignore_nulls)


class Max(AggregateFn):
    kind = KIND_MAX

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        super().__init__(_named("max", on, alias_name), on, ignore_nulls)


class Mean(AggregateFn):
    kind = KIND_MEAN

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        super().__init__(_named("mean", on, alias_name), on, ignore_nulls)


class Std(AggregateFn):
    """aggregate.py:250-326: Welford / Chan accumulator (M2, mean, count); here M2 is
    accumulated in a second pass around the group mean and partials are combined with
    the same parallel formula."""

    kind = KIND_STD

    def __init__(self, on: Optional[str] = None, ddof: int = 1, ignore_nulls: bool = True,
                 alias_name: Optional[str] = None):
        super().__init__(_named("std", on, alias_name), on, ignore_nulls, ddof)


def _unsupported(name):
    class _Unsupported(AggregateFn):
        def __init__(self, *a, **k):
            raise NotImplementedError(
                f"{name} is row / list based in the reference (python/ray/data/aggregate.py:330-489) "
                "and is not offloaded to the B200 engine; use the reference path for it")

    _Unsupported.__name__ = name
    return _Unsupported


AbsMax, Quantile, Unique = _unsupported("AbsMax"), _unsupported("Quantile"), _unsupported("Unique")


def check_aggs(aggs: Sequence[AggregateFn]):
    for a in aggs:
        if not isinstance(a, AggregateFn) or a.kind < 0:
            raise NotImplementedError(
                f"aggregate {a!r} is not one of the built-in Count/Sum/Min/Max/Mean/Std; "
                "user-defined AggregateFn callbacks stay on the reference path")
        if a.kind != KIND_COUNT and not isinstance(a.on, str):
            raise ValueError(f"aggregate {a.name!r} needs a column name (`on=`)")


def munge_names(names: List[str]) -> List[str]:
    """Duplicate result names get ``_2``, ``_3`` ... (_internal/table_block.py:171,
    arrow_block.py:515-522)."""
    seen: Dict[str, int] = {}
    out = []
    for n in names:
        c = seen.get(n, 0)
        out.append(n if c == 0 else f"{n}_{c + 1}")
        seen[n] = c + 1
    return out


def prune_columns(table: pa.Table, key_cols: List[str], aggs: Sequence[AggregateFn]) -> pa.Table:
    """SortAggregateTaskSpec._prune_unused_columns (aggregate_task_spec.py:76-109)."""
    cols = list(dict.fromkeys(list(key_cols) + [a.on for a in aggs if a.kind != KIND_COUNT]))
    if table.num_rows > 0 and table.num_columns > 0:
        return table.select(cols)
    return table


# ---------------------------------------------------------------------------
# device side
# ---------------------------------------------------------------------------

def _is_int_type(vt: int) -> bool:
    return vt not in (_cabi.F32, _cabi.F64)


def _next_pow2(x: int) -> int:
    p = 1
    while p < x:
        p <<= 1
    return p


DENSE_MAX_RANGE = 1 << 28  # XCHG_DENSE_MAX_RANGE

_KEY_RANGE = {_cabi.I8: 1 << 8, _cabi.U8: 1 << 8, _cabi.I16: 1 << 16, _cabi.U16: 1 << 16}


def _resolve_nulls(col: DeviceColumn) -> int:
    nc = col.null_count
    return int(nc.item()) if isinstance(nc, torch.Tensor) else int(nc)


def group_rows(rt: dev.Runtime, key_col: DeviceColumn, n: int):
    """K8 front half: -> (row group ids uint32[n], sorted group radix keys, G, key_type).
    Group ids follow the OUTPUT order (key ascending, NaN, then the null group)."""
    if key_col.kind != "fixed":
        raise NotImplementedError(f"group key of type {key_col.type} is not supported by the B200 engine")
    kt = dev.key_type_of(key_col.type)
    norm = rt.empty(n, torch.int64)
    rt.call("xchg_normalize_keys", _vp(key_col.data), kt, 0, _vp(key_col.validity), 0, None, n,
            _vp(norm), rt.stream)
    if _is_int_type(kt) and os.environ.get("XCHG_DENSE_GROUPS", "1") != "0":
        # integer keys over a small range: direct addressing instead of a hash table
        scratch = rt.zeros(32)
        mm = (ctypes.c_uint64 * 3)()
        rt.call("xchg_key_minmax", _vp(norm), _vp(key_col.validity), 0, n, _vp(scratch), mm, rt.stream)
        kmin, kmax, nvalid = int(mm[0]), int(mm[1]), int(mm[2])
        rng = (kmax - kmin + 1) if nvalid else 1
        if rng <= min(DENSE_MAX_RANGE, max(1 << 20, 4 * n)):
            kmin = kmin if nvalid else 0
            wsb = ctypes.c_uint64()
            _cabi.check(_cabi.load().xchg_dense_groups_bytes(rng, ctypes.byref(wsb)))
            ws = rt.empty(wsb.value)
            rgid = rt.empty(n, torch.int32)
            gkeys = rt.empty(min(rng, n) + 2, torch.int64)
            g_out = ctypes.c_uint64()
            rt.call("xchg_dense_groups", _vp(norm), _vp(key_col.validity), 0, n, kmin, rng, _vp(ws), wsb.value,
                    _vp(rgid), _vp(gkeys), ctypes.byref(g_out), rt.stream, launches=5)
            g = int(g_out.value)
            return rgid, gkeys[:g], g, kt
    cap = max(64, _next_pow2(2 * min(n, _KEY_RANGE.get(kt, n))))
    ws_bytes = ctypes.c_uint64()
    _cabi.check(_cabi.load().xchg_hash_table_bytes(cap, ctypes.byref(ws_bytes)))
    ws = rt.empty(ws_bytes.value)
    max_groups = min(n, cap) + 2
    slot = rt.empty(n, torch.int32)
    gkeys = rt.empty(max_groups, torch.int64)
    gslots = rt.empty(max_groups, torch.int32)
    g_out = ctypes.c_uint64()
    rt.call("xchg_hash_groups", _vp(norm), _vp(key_col.validity), 0, n, cap, _vp(ws), ws_bytes.value,
            _vp(slot), _vp(gkeys), _vp(gslots), ctypes.byref(g_out), rt.stream, launches=3)
    g = int(g_out.value)
    # output order of the groups: stable radix sort of their keys (K2 + K3)
    sws_bytes = dev.sort_workspace_bytes(g)
    sws = rt.empty(sws_bytes)
    order = rt.empty(g, torch.int32)
    sorted_keys = rt.empty(g, torch.int64)
    rt.call("xchg_sort_column", _vp(gkeys), _cabi.U64, 0, None, 0, None, g, _vp(order), _vp(sorted_keys),
            _vp(sws), sws_bytes, rt.stream, launches=10)
    rgid = rt.empty(n, torch.int32)
    rt.call("xchg_group_ids", _vp(order), _vp(gslots), g, _vp(slot), n, cap, _vp(ws), _vp(rgid),
            rt.stream, launches=2)
    return rgid, sorted_keys, g, kt


def _bytes_to_validity(rt: dev.Runtime, valid_bytes: torch.Tensor, g: int):
    bm = rt.zeros(((g + 31) // 32) * 4)
    cnt = rt.zeros(1, torch.int64)
    rt.call("xchg_bitmap_pack", _vp(valid_bytes), g, _vp(bm), _vp(cnt), rt.stream)
    return bm, cnt


def _state_field(col_name: str, vt: int) -> pa.Field:
    return pa.field(STATE_PREFIX + col_name, pa.binary(8 * STATE_WORDS), metadata={_VT_META: str(vt).encode()})


def is_partial_schema(schema: pa.Schema) -> bool:
    return ROWS_COL in schema.names


def aggregate_table(rt: dev.Runtime, table: DeviceTable, key: Optional[str],
                    aggs: Sequence[AggregateFn], finalize: bool = True) -> Optional[DeviceTable]:
    """Group ``table`` by ``key`` (``None`` = one global group) and aggregate.

    ``table`` holds either raw rows or partial rows (``is_partial_schema``); the result
    is the final table (``finalize=True``) or a partial table, one row per group,
    sorted by key.  Returns ``None`` for an empty input."""
    check_aggs(aggs)
    n = table.num_rows
    if n == 0:
        return None
    partial_in = is_partial_schema(table.schema)
    rgid = None
    g = 1
    key_col_out = None
    key_field = None
    if key is not None:
        kc = table.column(key)
        rgid, sorted_keys, g, kt = group_rows(rt, kc, n)
        kvals = rt.empty(g, torch.int64)
        rt.call("xchg_group_keys_out", _vp(sorted_keys), kt, g, _vp(kvals), rt.stream)
        out_type = pa.int64() if _is_int_type(kt) else pa.float64()
        kvalid, knulls = None, 0
        if kc.validity is not None and _resolve_nulls(kc) > 0:
            vb = rt.empty(g)
            vb.fill_(1)
            vb[g - 1:g].fill_(0)  # the null group is the last one
            kvalid, knulls = _bytes_to_validity(rt, vb, g)
        key_col_out = DeviceColumn(out_type, g, "fixed", 8, kvals.view(torch.uint8), None, kvalid, knulls)
        key_field = pa.field(key, out_type)

    rows = rt.zeros(g, torch.int64)
    if partial_in:
        rt.call("xchg_agg_merge_rows", _vp(rgid), _vp(table.column(ROWS_COL).data), n, _vp(rows), rt.stream)
    else:
        rt.call("xchg_agg_count_rows", _vp(rgid), n, g, _vp(rows), rt.stream)

    # one statistics block per distinct value column
    value_cols = list(dict.fromkeys(a.on for a in aggs if a.kind != KIND_COUNT))
    stats: Dict[str, Tuple[torch.Tensor, Optional[torch.Tensor], int]] = {}
    for name in value_cols:
        need_m2 = (not finalize) or any(a.kind == KIND_STD and a.on == name for a in aggs)
        st = rt.empty(5 * g, torch.int64)
        m2 = rt.empty(g, torch.float64) if need_m2 else None
        rt.call("xchg_agg_stats_init", _vp(st), _vp(m2), g, rt.stream, launches=2)
        if partial_in:
            sc = table.column(STATE_PREFIX + name)
            fld = table.schema.field(STATE_PREFIX + name)
            vt = int(fld.metadata[_VT_META])
            rt.call("xchg_agg_merge_states", _vp(rgid), _vp(sc.data), n, int(_is_int_type(vt)), g, _vp(st),
                    _vp(m2), rt.stream, launches=2)
        else:
            vc = table.column(name)
            if vc.kind != "fixed":
                raise NotImplementedError(f"aggregate on column {name!r} of type {vc.type} is not supported")
            vt = dev.key_type_of(vc.type)
            # min / max / NaN counts only where something reads them: Min, Max, or a partial
            # row (whose consumer is unknown here)
            want_minmax = int((not finalize) or any(a.kind in (KIND_MIN, KIND_MAX) and a.on == name for a in aggs))
            rt.call("xchg_agg_accumulate", _vp(rgid), _vp(vc.data), vt, _vp(vc.validity), 0, n, g, _vp(st),
                    want_minmax, _vp(rows), rt.stream, launches=2)
            if need_m2:
                rt.call("xchg_agg_accumulate_m2", _vp(rgid), _vp(vc.data), vt, _vp(vc.validity), 0, n, g,
                        _vp(st), _vp(m2), rt.stream)
        stats[name] = (st, m2, vt)

    cols: List[DeviceColumn] = []
    fields: List[pa.Field] = []
    if key_col_out is not None:
        cols.append(key_col_out)
        fields.append(key_field)
    if not finalize:
        cols.append(DeviceColumn(pa.int64(), g, "fixed", 8, rows.view(torch.uint8)))
        fields.append(pa.field(ROWS_COL, pa.int64()))
        for name in value_cols:
            st, m2, vt = stats[name]
            packed = rt.empty(g * STATE_WORDS, torch.int64)
            rt.call("xchg_agg_pack_states", g, _vp(st), _vp(m2), _vp(packed), rt.stream)
            cols.append(DeviceColumn(pa.binary(8 * STATE_WORDS), g, "fixed", 8 * STATE_WORDS,
                                     packed.view(torch.uint8)))
            fields.append(_state_field(name, vt))
        return DeviceTable([f.name for f in fields], cols, g, pa.schema(fields))

    names = munge_names(([key] if key is not None else []) + [a.name for a in aggs])
    names = names[1:] if key is not None else names
    for a, out_name in zip(aggs, names):
        out = rt.empty(g, torch.int64)
        vb = rt.empty(g)
        if a.kind == KIND_COUNT:
            rt.call("xchg_agg_finalize", KIND_COUNT, 0, 1, _cabi.I64, g, _vp(rows), None, None, _vp(out),
                    _vp(vb), rt.stream)
            typ = pa.int64()
        else:
            st, m2, vt = stats[a.on]
            rt.call("xchg_agg_finalize", a.kind, a.ddof, int(a.ignore_nulls), vt, g, _vp(rows), _vp(st),
                    _vp(m2), _vp(out), _vp(vb), rt.stream)
            int_out = a.kind in (KIND_SUM, KIND_MIN, KIND_MAX) and _is_int_type(vt)
            typ = pa.int64() if int_out else pa.float64()
        validity, nulls = (None, 0) if a.kind == KIND_COUNT else _bytes_to_validity(rt, vb, g)
        cols.append(DeviceColumn(typ, g, "fixed", 8, out.view(torch.uint8), None, validity, nulls))
        fields.append(pa.field(out_name, typ))
    return DeviceTable([f.name for f in fields], cols, g, pa.schema(fields))


def group_cut_points(rt: dev.Runtime, result: DeviceTable, key: str, boundaries: Sequence[tuple]) -> List[int]:
    """Range-partition cut points of an aggregate result / partial table (sorted by
    key ascending) under the reference's boundary rule (K4): partition i = groups
    ``[cut[i-1], cut[i])`` -- what ``sort_and_partition`` + per-partition ``combine``
    produce in SortAggregateTaskSpec.map (aggregate_task_spec.py:38-59)."""
    from . import keys as keymod

    if not boundaries:
        return []
    kc = result.column(key)
    g = result.num_rows
    kt = dev.key_type_of(kc.type)
    norm = rt.empty(g, torch.int64)
    rt.call("xchg_normalize_keys", _vp(kc.data), kt, 0, _vp(kc.validity), 0, None, g, _vp(norm), rt.stream)
    lower, upper, flags = keymod.normalize_boundaries(boundaries, [kt], [False])
    thr = keymod.last_class_threshold(kt, _resolve_nulls(kc))
    return dev.partition_cuts(rt, [norm], g, lower, upper, flags, [False], [thr[0]], [thr[1]])

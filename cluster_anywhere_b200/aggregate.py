"""Aggregations of ``groupby().aggregate()`` and their device implementation.

Host surface: ``Count / Sum / Min / Max / Mean / Std / AbsMax`` with the reference's
constructor arguments, result names and null policy (python/ray/data/aggregate.py:
97-354, _internal/null_aggregate.py:11-25).  ``Quantile / Unique`` are row /
list based in the reference (:367-489) and stay on the reference path: asking this
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

KIND_COUNT, KIND_SUM, KIND_MIN, KIND_MAX, KIND_MEAN, KIND_STD, KIND_ABSMAX = range(7)
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


class AbsMax(AggregateFn):
    """aggregate.py:330-354: ``max(0, |x|)`` over the group's non-null values (a NaN never wins
    Python's ``max(a, abs(r))``).  Computed from the group's min and max at finalisation."""

    kind = KIND_ABSMAX

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        super().__init__(_named("abs_max", on, alias_name), on, ignore_nulls)


def _unsupported(name):
    class _Unsupported(AggregateFn):
        def __init__(self, *a, **k):
            raise NotImplementedError(
                f"{name} is row / list based in the reference (python/ray/data/aggregate.py:367-489) "
                "and is not offloaded to the B200 engine; use the reference path for it")

    _Unsupported.__name__ = name
    return _Unsupported


Quantile, Unique = _unsupported("Quantile"), _unsupported("Unique")


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


def group_rows(rt: dev.Runtime, key_col: DeviceColumn, n: int, want_rows: bool = False):
    """K8 front half: -> (row group ids uint32[n], sorted group radix keys, G, key_type,
    rows per group or None).  Group ids follow the OUTPUT order (key ascending, NaN, then
    the null group).  ``want_rows``: the direct-address path counts the rows of every group
    in its marking pass (the hash path returns None: the caller counts them)."""
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
            grows = rt.empty(min(rng, n) + 2, torch.int64) if want_rows else None
            g_out = ctypes.c_uint64()
            rt.call("xchg_dense_groups", _vp(norm), _vp(key_col.validity), 0, n, kmin, rng, _vp(ws), wsb.value,
                    _vp(rgid), _vp(gkeys), _vp(grows), ctypes.byref(g_out), rt.stream, launches=6)
            g = int(g_out.value)
            return rgid, gkeys[:g], g, kt, (grows[:g] if want_rows else None)
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
    return rgid, sorted_keys, g, kt, None


def group_rows_sorted(rt: dev.Runtime, table: DeviceTable, key_names: Sequence[str]):
    """General grouping (several key columns, string / binary keys): stable sort of the
    rows by the keys, per-column change marks on the sorted order, scan of the marks.
    -> (row group ids uint32[n] in OUTPUT order, G, original row id of every group's
    first row uint32[G])."""
    n = table.num_rows
    perm, norms = dev.sort_permutation(rt, table, list(key_names), [False] * len(key_names))
    flags = rt.zeros(n)
    for ci, name in enumerate(key_names):
        col = table.column(name)
        if dev.is_string_key(col):
            rt.call("xchg_group_flags_bytes", _vp(dev._string_offsets(rt, col)), _vp(col.data),
                    _vp(col.validity), 0, _vp(perm), n, _vp(flags), rt.stream)
            continue
        if col.kind != "fixed":
            raise NotImplementedError(f"group key of type {col.type} is not supported by the B200 engine")
        nk = norms[0] if ci == 0 and norms and norms[0] is not None else None
        if nk is None:
            nk = rt.empty(n, torch.int64)
            rt.call("xchg_normalize_keys", _vp(col.data), dev.key_type_of(col.type), 0, _vp(col.validity), 0,
                    _vp(perm), n, _vp(nk), rt.stream)
        rt.call("xchg_group_flags_fixed", _vp(nk), _vp(col.validity), 0, _vp(perm), n, _vp(flags), rt.stream)
    wsb = ctypes.c_uint64()
    _cabi.check(_cabi.load().xchg_groups_from_flags_bytes(n, ctypes.byref(wsb)))
    ws = rt.empty(wsb.value)
    rgid = rt.empty(n, torch.int32)
    first = rt.empty(n, torch.int32)
    g_out = ctypes.c_uint64()
    rt.call("xchg_groups_from_flags", _vp(flags), _vp(perm), n, _vp(ws), wsb.value, _vp(rgid), _vp(first),
            ctypes.byref(g_out), rt.stream, launches=3)
    g = int(g_out.value)
    return rgid, g, first[:g]


def _key_columns_out(rt: dev.Runtime, table: DeviceTable, key_names: Sequence[str], first: torch.Tensor, g: int):
    """Key columns of the result: the key values of every group's first row.  Integer
    keys come out as int64 and float keys as double, as the reference rebuilds them from
    Python scalars; string / binary keys keep their type."""
    cols, fields = [], []
    for name in key_names:
        kc = table.column(name)
        if dev.is_string_key(kc):
            cols.append(dev.take_column(rt, kc, first, g))
            fields.append(pa.field(name, kc.type))
            continue
        kt = dev.key_type_of(kc.type)
        nk = rt.empty(g, torch.int64)
        rt.call("xchg_normalize_keys", _vp(kc.data), kt, 0, _vp(kc.validity), 0, _vp(first), g, _vp(nk), rt.stream)
        kvals = rt.empty(g, torch.int64)
        rt.call("xchg_group_keys_out", _vp(nk), kt, g, _vp(kvals), rt.stream)
        out_type = pa.int64() if _is_int_type(kt) else pa.float64()
        kvalid, knulls = None, 0
        if kc.validity is not None and _resolve_nulls(kc) > 0:
            kvalid = rt.zeros(((g + 31) // 32) * 4)
            cnt = rt.zeros(1, torch.int64)
            rt.call("xchg_gather_bitmap", _vp(kc.validity), 0, _vp(first), g, _vp(kvalid), _vp(cnt), rt.stream)
            knulls = cnt
        cols.append(DeviceColumn(out_type, g, "fixed", 8, kvals.view(torch.uint8), None, kvalid, knulls))
        fields.append(pa.field(name, out_type))
    return cols, fields


def null_typed_like_reference(t: pa.Table) -> pa.Table:
    """A FINAL aggregate block is rebuilt by the reference row by row from Python scalars
    (ArrowBlockBuilder over the rows of aggregate_combined_blocks, arrow_block.py:593-660),
    so a column whose every value is None comes out with Arrow type ``null`` (e.g. the key
    column of the block that holds only the null group), and a block without groups is the
    schema-less empty table (arrow_block.py:316).  Same here, on the host copy."""
    if t.num_rows == 0:
        return pa.table({})
    for i, name in enumerate(t.schema.names):
        col = t.column(i)
        if col.null_count == t.num_rows and not pa.types.is_null(col.type):
            t = t.set_column(i, name, pa.nulls(t.num_rows))
    return t


def key_list(key) -> List[str]:
    """``None`` / one column name / a list of names -> list of key columns."""
    if key is None:
        return []
    if isinstance(key, str):
        return [key]
    return list(key)


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
    keys = key_list(key)
    key_cols_out: List[DeviceColumn] = []
    key_fields: List[pa.Field] = []
    counted = None
    general = len(keys) > 1 or (len(keys) == 1 and table.column(keys[0]).kind != "fixed")
    if general:
        rgid, g, first = group_rows_sorted(rt, table, keys)
        key_cols_out, key_fields = _key_columns_out(rt, table, keys, first, g)
    elif keys:
        key = keys[0]
        kc = table.column(key)
        rgid, sorted_keys, g, kt, counted = group_rows(rt, kc, n, want_rows=not partial_in)
        kvals = rt.empty(g, torch.int64)
        rt.call("xchg_group_keys_out", _vp(sorted_keys), kt, g, _vp(kvals), rt.stream)
        out_type = pa.int64() if _is_int_type(kt) else pa.float64()
        kvalid, knulls = None, 0
        if kc.validity is not None and _resolve_nulls(kc) > 0:
            vb = rt.empty(g)
            vb.fill_(1)
            vb[g - 1:g].fill_(0)  # the null group is the last one
            kvalid, knulls = _bytes_to_validity(rt, vb, g)
        key_cols_out = [DeviceColumn(out_type, g, "fixed", 8, kvals.view(torch.uint8), None, kvalid, knulls)]
        key_fields = [pa.field(key, out_type)]

    rows = counted if counted is not None else rt.zeros(g, torch.int64)
    if counted is not None:
        pass  # the direct-address grouping counted the rows of every group on its way
    elif partial_in:
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
            want_minmax = int((not finalize)
                              or any(a.kind in (KIND_MIN, KIND_MAX, KIND_ABSMAX) and a.on == name for a in aggs))
            rt.call("xchg_agg_accumulate", _vp(rgid), _vp(vc.data), vt, _vp(vc.validity), 0, n, g, _vp(st),
                    want_minmax, _vp(rows), rt.stream, launches=2)
            if need_m2:
                rt.call("xchg_agg_accumulate_m2", _vp(rgid), _vp(vc.data), vt, _vp(vc.validity), 0, n, g,
                        _vp(st), _vp(m2), rt.stream)
        stats[name] = (st, m2, vt)

    cols: List[DeviceColumn] = []
    fields: List[pa.Field] = []
    cols.extend(key_cols_out)
    fields.extend(key_fields)
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

    names = munge_names(keys + [a.name for a in aggs])[len(keys):]
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
            int_out = a.kind in (KIND_SUM, KIND_MIN, KIND_MAX, KIND_ABSMAX) and _is_int_type(vt)
            typ = pa.int64() if int_out else pa.float64()
        validity, nulls = (None, 0) if a.kind == KIND_COUNT else _bytes_to_validity(rt, vb, g)
        cols.append(DeviceColumn(typ, g, "fixed", 8, out.view(torch.uint8), None, validity, nulls))
        fields.append(pa.field(out_name, typ))
    return DeviceTable([f.name for f in fields], cols, g, pa.schema(fields))


def _cuts_by_merged_sort(rt: dev.Runtime, result: DeviceTable, cols: List[str],
                         boundaries: Sequence[tuple]) -> List[int]:
    """Cut points for key tuples that mix string and numeric columns.  ``result`` holds one
    row per DISTINCT key tuple, sorted ascending; the cut of a boundary is the number of
    groups strictly before it (find_partition_index, _internal/util.py:728-803, all columns
    ascending: the left end of the narrowing search; a None component sorts after every
    value, as NULL_SENTINEL does).  The boundary tuples are appended to the group keys and
    the whole is grouped once more (stable sort + run marks): the groups before a boundary
    are the runs before its run that contain a data row."""
    G, nb = result.num_rows, len(boundaries)
    # A None component on a NUMERIC column ends the reference's search after every row that
    # shares the components before it: pyarrow hands numpy a float array with NaN for the
    # nulls, `col_vals == None` matches nothing, and NULL_SENTINEL compares above everything
    # (util.py:761-770, 800-801), so left == right == end of the run and the loop returns.
    # On a string column the nulls ARE replaced by the sentinel and the search goes on inside
    # the null rows.  Hence: from the first numeric None on, the boundary is "after every row
    # with this prefix": nulls in the remaining components and an UPPER bound.
    is_str = [dev.is_string_key(result.column(c)) for c in cols]
    tuples, upper = [], []
    for bnd in boundaries:
        bnd = tuple(bnd) + (None,) * (len(cols) - len(bnd))
        cut_at = next((i for i, v in enumerate(bnd) if v is None and not is_str[i]), None)
        if cut_at is not None:
            bnd = bnd[:cut_at] + (None,) * (len(cols) - cut_at)
        tuples.append(bnd)
        upper.append(cut_at is not None)
    try:
        btab = pa.table({c: pa.array([t[i] for t in tuples], type=result.column(c).type)
                         for i, c in enumerate(cols)})
    except (pa.ArrowInvalid, pa.ArrowTypeError) as e:
        raise NotImplementedError(f"boundaries {list(boundaries)!r} do not convert to the key types: {e}")
    bdev = dev.stage_table(rt, btab)
    merged_cols = [dev.concat_columns(rt, result.column(c), bdev.column(c)) for c in cols]
    merged = DeviceTable(list(cols), merged_cols, G + nb, pa.schema([(c, result.column(c).type) for c in cols]))
    rgid, g, first = group_rows_sorted(rt, merged, cols)
    gid_b = rgid[G:G + nb].cpu().numpy().view(np.uint32).astype(np.int64)
    first_h = first.cpu().numpy().view(np.uint32).astype(np.int64)
    boundary_only = sorted(set(int(x) for x in gid_b if first_h[int(x)] >= G))
    cuts = []
    for x, up in zip(gid_b, upper):
        x = int(x)
        before = x - sum(1 for y in boundary_only if y < x)  # runs with a data row before this run
        has_data_row = first_h[x] < G                         # data rows come first inside a run
        cuts.append(before + (1 if (up and has_data_row) else 0))
    return cuts


def group_cut_points(rt: dev.Runtime, result: DeviceTable, key, boundaries: Sequence[tuple]) -> List[int]:
    """Range-partition cut points of an aggregate result / partial table (sorted by
    key ascending) under the reference's boundary rule (K4): partition i = groups
    ``[cut[i-1], cut[i])`` -- what ``sort_and_partition`` + per-partition ``combine``
    produce in SortAggregateTaskSpec.map (aggregate_task_spec.py:38-59).  ``key`` = one
    column name or a list; a string key is searched on the (sorted) result column itself."""
    from . import keys as keymod

    if not boundaries:
        return []
    cols = key_list(key)
    g = result.num_rows
    if any(dev.is_string_key(result.column(c)) for c in cols):
        if len(cols) != 1:
            return _cuts_by_merged_sort(rt, result, cols, boundaries)
        return dev.string_partition_cuts(rt, result.column(cols[0]), boundaries, False)
    norms, ktypes, thr = [], [], []
    for c in cols:
        kc = result.column(c)
        kt = dev.key_type_of(kc.type)
        norm = rt.empty(g, torch.int64)
        rt.call("xchg_normalize_keys", _vp(kc.data), kt, 0, _vp(kc.validity), 0, None, g, _vp(norm), rt.stream)
        norms.append(norm)
        ktypes.append(kt)
        thr.append(keymod.last_class_threshold(kt, _resolve_nulls(kc)))
    desc = [False] * len(cols)
    lower, upper, flags = keymod.normalize_boundaries(boundaries, ktypes, desc)
    return dev.partition_cuts(rt, norms, g, lower, upper, flags, desc, [t[0] for t in thr], [t[1] for t in thr])


# ---------------------------------------------------------------------------
# Group boundaries of a key-sorted block, distinct values
# ---------------------------------------------------------------------------

def block_boundaries(rt: dev.Runtime, table: DeviceTable, key_names: Sequence[str]) -> np.ndarray:
    """Boundaries of the groups of a block that is SORTED by ``key_names``: the start index
    of every group and ``num_rows`` at the end -- ``_get_block_boundaries``
    (python/ray/data/block.py:515-592), which GroupedData.map_groups applies to every block
    of the sorted dataset (grouped_data.py:196-213).  NaN == NaN and null == null, per column.
    Run marks per key column in storage order (no permutation) + scan, on the device."""
    n = table.num_rows
    if n == 0:
        return np.array([0, 0], dtype=int)
    if not key_names:
        return np.array([0, n], dtype=int)
    flags = rt.zeros(n)
    for name in key_names:
        col = table.column(name)
        if dev.is_string_key(col):
            rt.call("xchg_group_flags_bytes", _vp(dev._string_offsets(rt, col)), _vp(col.data),
                    _vp(col.validity), 0, None, n, _vp(flags), rt.stream)
            continue
        if col.kind != "fixed":
            raise NotImplementedError(f"group key of type {col.type} is not supported by the B200 engine")
        nk = rt.empty(n, torch.int64)
        rt.call("xchg_normalize_keys", _vp(col.data), dev.key_type_of(col.type), 0, _vp(col.validity), 0, None, n,
                _vp(nk), rt.stream)
        rt.call("xchg_group_flags_fixed", _vp(nk), _vp(col.validity), 0, None, n, _vp(flags), rt.stream)
    wsb = ctypes.c_uint64()
    _cabi.check(_cabi.load().xchg_groups_from_flags_bytes(n, ctypes.byref(wsb)))
    ws = rt.empty(wsb.value)
    rgid = rt.empty(n, torch.int32)
    first = rt.empty(n, torch.int32)
    g_out = ctypes.c_uint64()
    rt.call("xchg_groups_from_flags", _vp(flags), None, n, _vp(ws), wsb.value, _vp(rgid), _vp(first),
            ctypes.byref(g_out), rt.stream, launches=3)
    g = int(g_out.value)
    starts = first[:g].cpu().numpy().view(np.uint32).astype(int)
    return np.concatenate([starts, [n]]).astype(int)


def _get_block_boundaries(columns: Sequence[np.ndarray]) -> np.ndarray:
    """The reference's entry point (block.py:515): key columns of a sorted block as 1-D numpy
    arrays (object arrays with None for strings, float arrays with NaN) -> boundaries."""
    arrays = {}
    for i, arr in enumerate(columns):
        arr = np.asarray(arr)
        if arr.dtype == object:
            arrays[f"c{i}"] = pa.array(arr.tolist())
        else:
            arrays[f"c{i}"] = pa.array(arr)
    table = pa.table(arrays)
    rt = dev.default_runtime()
    return block_boundaries(rt, dev.stage_table(rt, table), list(arrays))


def unique_values(rt: dev.Runtime, table: DeviceTable, column: str) -> DeviceColumn:
    """Distinct values of one column in ascending order, null last when present
    (``Dataset.unique`` -> ``Unique`` aggregate, python/ray/data/dataset.py:2182-2221,
    aggregate.py:445-489: ``pac.unique`` per block, set union across blocks)."""
    n = table.num_rows
    kc = table.column(column)
    if n == 0:
        return kc
    if kc.kind == "fixed":
        rgid, sorted_keys, g, kt, _ = group_rows(rt, kc, n)
        # first row of every group: any row will do, the values are equal
        first = rt.empty(g, torch.int32)
        rt.call("xchg_group_first_rows", _vp(rgid), n, g, _vp(first), rt.stream)
    else:
        rgid, g, first = group_rows_sorted(rt, table, [column])
    return dev.take_column(rt, kc, first, g)

"""Host-side mirror of the device key normalisation (csrc/radix_sort.cu) for the
handful of *boundary* values a partition search needs.

Rows are normalised on the GPU (K2).  Boundaries -- at most a few thousand tuples
-- are normalised here with the same order-preserving transform, and turned into
the ``lower`` / ``upper`` / ``flags`` triples ``xchg_partition_bounds`` takes.

The transform (unsigned 64-bit radix key, always sorted ascending):
  signed int W bits   u = x + 2^(W-1);          descending: 2^W-1-u
  unsigned            u = x;                    descending: 2^W-1-u
  float               IEEE bits, sign-flipped (-0.0 == +0.0); descending: ~u
  NaN                 2^W-2  (after every number, both directions)
  null                one past the value range (2^W for W<64, all-ones otherwise)

Semantics restated from find_partition_index
(python/ray/data/_internal/util.py:728-803): ``lo`` counts keys strictly before
the boundary in the column's direction, ``hi`` counts keys before-or-equal; a
boundary that no stored value can equal (2.5 on an int column, 0.1 on a float32
column, out-of-range ints) has ``hi == lo`` (flag bit 0); ``None`` sorts after
everything (flag bit 2).
"""
from __future__ import annotations

import math
from typing import Sequence, Tuple

import numpy as np
import pyarrow as pa

from . import _cabi

FLAG_EMPTY = 1
FLAG_STOP = 2
FLAG_AFTER_ALL = 4
FLAG_NAN = 8

_INT_INFO = {
    _cabi.I8: (8, True), _cabi.I16: (16, True), _cabi.I32: (32, True), _cabi.I64: (64, True),
    _cabi.U8: (8, False), _cabi.U16: (16, False), _cabi.U32: (32, False), _cabi.U64: (64, False),
}
_U64 = (1 << 64) - 1


def _norm_int(x: int, bits: int, signed: bool, desc: bool) -> int:
    u = x + (1 << (bits - 1)) if signed else x
    return ((1 << bits) - 1 - u) if desc else u


def _norm_float(x: float, bits: int, desc: bool) -> int:
    """x must already be representable in the column's float type."""
    if x != x:
        return (1 << bits) - 2
    if bits == 32:
        b = int(np.float32(0.0 if x == 0 else x).view(np.uint32))
    else:
        b = int(np.float64(0.0 if x == 0 else x).view(np.uint64))
    sign = 1 << (bits - 1)
    mask = (1 << bits) - 1
    t = (~b & mask) if (b & sign) else (b | sign)
    return (~t & mask) if desc else t


def _null_key(bits: int) -> int:
    return _U64 if bits == 64 else (1 << bits)


def _float_neighbours(v, bits: int) -> Tuple[float, float]:
    """(largest representable <= v, smallest representable >= v) in the column type.
    ``v`` is a Python int or float; comparisons are exact (Python compares
    int/float exactly)."""
    ft = np.float32 if bits == 32 else np.float64
    with np.errstate(over="ignore"):
        c = ft(v)
    cf = float(c)
    if cf == v:
        return cf, cf
    if cf < v:
        up = float(np.nextafter(c, ft(np.inf)))
        return cf, up
    dn = float(np.nextafter(c, ft(-np.inf)))
    return dn, cf


def normalize_boundary_value(value, key_type: int, desc: bool):
    """-> (lower, upper, flags) for one component of a boundary tuple."""
    if value is None:
        return _U64, _U64, FLAG_AFTER_ALL
    if isinstance(value, (np.generic,)):
        value = value.item()
    if isinstance(value, bool):
        value = int(value)
    if key_type in _INT_INFO:
        bits, signed = _INT_INFO[key_type]
        lo_v = -(1 << (bits - 1)) if signed else 0
        hi_v = (1 << (bits - 1)) - 1 if signed else (1 << bits) - 1
        if isinstance(value, float) and value != value:
            # NaN boundary: numpy orders NaN after every number; null rows reach
            # numpy as NaN (int column with nulls -> float array), i.e. as equals
            k = _null_key(bits)
            return k, k, FLAG_NAN
        exact = isinstance(value, int) or float(value).is_integer()
        if isinstance(value, float) and math.isinf(value):
            exact = False
        if not desc:
            # lo = #(x < v) = #(x < ceil(v))
            if isinstance(value, float) and math.isinf(value):
                c = hi_v + 1 if value > 0 else lo_v - 1
            else:
                c = value if isinstance(value, int) else math.ceil(value)
            if c > hi_v:
                k = _null_key(bits)
                return k, k, FLAG_EMPTY
            if c < lo_v:
                return 0, 0, FLAG_EMPTY
            k = _norm_int(c, bits, signed, False)
            return k, k, (0 if exact else FLAG_EMPTY)
        # descending: lo = #(x > v) = #(nk(x) < nk(floor(v)))
        if isinstance(value, float) and math.isinf(value):
            f = hi_v + 1 if value > 0 else lo_v - 1
        else:
            f = value if isinstance(value, int) else math.floor(value)
        if f > hi_v:
            return 0, 0, FLAG_EMPTY
        if f < lo_v:
            k = _null_key(bits)
            return k, k, FLAG_EMPTY
        k = _norm_int(f, bits, signed, True)
        return k, k, (0 if exact else FLAG_EMPTY)
    if key_type in (_cabi.F32, _cabi.F64):
        bits = 32 if key_type == _cabi.F32 else 64
        if isinstance(value, float) and value != value:
            # equal range = NaN rows and null rows (both are NaN to numpy)
            return _norm_float(value, bits, desc), _null_key(bits), FLAG_NAN
        dn, up = _float_neighbours(value, bits)
        exact = dn == up
        pick = dn if desc else up
        k = _norm_float(pick, bits, desc)
        return k, k, (0 if exact else FLAG_EMPTY)
    raise NotImplementedError(f"boundary normalisation for key type {key_type}")


def normalize_boundaries(boundaries: Sequence[tuple], key_types: Sequence[int],
                         descending: Sequence[bool]):
    """Boundary tuples -> (lower[nb,ncols], upper[nb,ncols], flags[nb,ncols]).  A key type of
    ``None`` marks a string / binary column: its components travel as bytes
    (``pack_string_components``); here they only get the STOP / None flags."""
    nb, ncols = len(boundaries), len(key_types)
    lower = np.zeros((nb, ncols), dtype=np.uint64)
    upper = np.zeros((nb, ncols), dtype=np.uint64)
    flags = np.zeros((nb, ncols), dtype=np.uint8)
    for i, b in enumerate(boundaries):
        if not isinstance(b, (tuple, list)):
            b = (b,)
        if len(b) > ncols:
            raise ValueError(f"boundary {b!r} has more components than sort keys ({ncols})")
        for c in range(ncols):
            if c >= len(b):
                flags[i, c] = FLAG_STOP
                continue
            if key_types[c] is None:
                flags[i, c] = FLAG_AFTER_ALL if b[c] is None else 0
                continue
            lo, up, fl = normalize_boundary_value(b[c], key_types[c], bool(descending[c]))
            lower[i, c], upper[i, c], flags[i, c] = lo, up, fl
    return lower, upper, flags


def pack_string_components(boundaries: Sequence[tuple], c: int):
    """Component ``c`` of every boundary tuple as an Arrow-binary-like pair
    (offsets int64[nb+1], bytes uint8[]); missing and None components are empty."""
    vals = []
    for b in boundaries:
        if not isinstance(b, (tuple, list)):
            b = (b,)
        v = b[c] if c < len(b) else None
        if v is None:
            vals.append(b"")
        elif isinstance(v, str):
            vals.append(v.encode("utf-8"))
        elif isinstance(v, (bytes, bytearray, memoryview)):
            vals.append(bytes(v))
        else:
            raise TypeError(f"boundary component {v!r} does not compare with a string key column")
    off = np.zeros(len(vals) + 1, dtype=np.int64)
    if vals:
        off[1:] = np.cumsum([len(v) for v in vals])
    blob = np.frombuffer(b"".join(vals), dtype=np.uint8).copy() if off[-1] else np.zeros(0, np.uint8)
    return off, blob


def last_class_threshold(key_type: int, null_count: int):
    """-> (threshold, has_last): normalised keys >= threshold are NaN / null rows,
    the values numpy sees as NaN in ``table[col].to_numpy()`` (util.py:761)."""
    if key_type in (_cabi.F32, _cabi.F64):
        bits = 32 if key_type == _cabi.F32 else 64
        return (1 << bits) - 2, True
    bits, _ = _INT_INFO[key_type]
    return _null_key(bits), null_count > 0

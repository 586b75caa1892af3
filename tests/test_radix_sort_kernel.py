"""xchg_sort_column called directly through the C-ABI: the hybrid (MSD levels + local
shared-memory sort) against numpy's stable argsort of the same radix keys and against
the all-global LSD passes, on key distributions that reach every branch -- segments of
every local-sort size class, segments that stay large for several levels, a run of
equal keys that never becomes small, null rows on 64-bit integers (validity place),
descending order, a chained second sort (perm_in)."""
import ctypes

import numpy as np
import pytest
import torch

from cluster_anywhere_b200 import _cabi
from cluster_anywhere_b200 import device as dev

pytestmark = pytest.mark.gpu


def _sort(rt, keys_np, key_type, desc=False, validity=None, perm_in=None, algorithm=0):
    n = len(keys_np)
    rt.call("xchg_set_sort_algorithm", algorithm, launches=0)
    try:
        kb = rt.empty(max(keys_np.nbytes, 8))
        dev._h2d(kb[: keys_np.nbytes], keys_np.view(np.uint8))
        vb = None
        if validity is not None:
            packed = np.packbits(validity, bitorder="little")
            vb = rt.zeros(len(packed) + 8)
            dev._h2d(vb[: len(packed)], packed)
        pin = None
        if perm_in is not None:
            pin = rt.empty(n, torch.int32)
            dev._h2d(pin.view(torch.uint8), perm_in.astype(np.uint32).view(np.uint8))
        ws_bytes = dev.sort_workspace_bytes(n)
        ws = rt.empty(ws_bytes)
        perm = rt.empty(n, torch.int32)
        norm = rt.empty(n, torch.int64)
        rt.call("xchg_sort_column", dev._vp(kb), key_type, int(desc), dev._vp(vb) if vb is not None else None, 0,
                dev._vp(pin) if pin is not None else None, n, dev._vp(perm), dev._vp(norm), dev._vp(ws),
                ws_bytes, rt.stream, launches=0)
        rt.synchronize()
        return perm.cpu().numpy().view(np.uint32), norm.cpu().numpy().view(np.uint64)
    finally:
        rt.call("xchg_set_sort_algorithm", 0, launches=0)


def _check_u64(rt, keys, perm_in=None):
    exp_src = keys if perm_in is None else keys[perm_in]
    order = np.argsort(exp_src, kind="stable")
    exp = order.astype(np.uint32) if perm_in is None else perm_in[order].astype(np.uint32)
    got, norm = _sort(rt, keys, _cabi.U64, perm_in=perm_in)
    assert np.array_equal(norm, exp_src[order])
    assert np.array_equal(got, exp)
    lsd, norm_l = _sort(rt, keys, _cabi.U64, perm_in=perm_in, algorithm=1)
    assert np.array_equal(lsd, exp) and np.array_equal(norm_l, norm)


def test_uniform_random_two_global_levels():
    rt = dev.default_runtime()
    rng = np.random.default_rng(1)
    keys = rng.integers(0, 2**64, 3_000_017, dtype=np.uint64)
    _check_u64(rt, keys)


def test_parents_just_above_the_local_capacity_group_their_children():
    """256 segments of ~11 700 rows (just above the 8192-row local capacity): each splits
    into 256 children of ~46 rows, which must be packed into a few local-sort jobs, not
    sorted one CTA each."""
    rt = dev.default_runtime()
    rng = np.random.default_rng(3)
    n = 3_000_000
    keys = (rng.integers(0, 256, n, dtype=np.uint64) << np.uint64(56)) | rng.integers(0, 2**48, n, dtype=np.uint64)
    _check_u64(rt, keys)


@pytest.mark.parametrize("places", [3, 4, 5, 8])
def test_active_places(places):
    rt = dev.default_runtime()
    rng = np.random.default_rng(places)
    keys = rng.integers(0, 2**(8 * places), 400_003, dtype=np.uint64) if places < 8 else \
        rng.integers(0, 2**64, 400_003, dtype=np.uint64)
    # spread the active bytes over non-adjacent places
    if places == 3:
        keys = (keys & 0xFF) | ((keys & 0xFF00) << 16) | ((keys & 0xFF0000) << 32)
    _check_u64(rt, keys)


def test_every_size_class_and_deep_segments():
    rt = dev.default_runtime()
    rng = np.random.default_rng(7)
    parts = []
    # top byte t selects a child of the first level with a chosen size
    for t, size in enumerate([1, 37, 300, 512, 513, 800, 1024, 1025, 1500, 2048, 2049, 3000, 4096, 4097,
                              5000, 8192, 8193, 9000, 70_000]):
        low = rng.integers(0, 2**56, size, dtype=np.uint64)
        parts.append((np.uint64(t) << np.uint64(56)) | low)
    # a segment whose rows agree in the top FIVE places: stays large for five levels
    deep = rng.integers(0, 2**24, 20_000, dtype=np.uint64)
    parts.append(np.uint64(0xAB12345678000000) | deep)
    # 30 000 equal keys: large until the places run out
    parts.append(np.full(30_000, 0xCD00000000000011, dtype=np.uint64))
    # duplicates-heavy segment: 50 distinct values
    parts.append(np.uint64(0xEE00000000000000) | rng.integers(0, 50, 25_000, dtype=np.uint64))
    # local-sort corner cases (segments of <= 4096 rows with >= 3 places left):
    # one long out-of-order run (equal in the two top remaining places) -> plain LSD passes
    parts.append(np.uint64(0x7700000000000000) | rng.integers(0, 2**40, 3000, dtype=np.uint64))
    # many short runs (500 distinct values of the two top remaining places, ~4 rows each)
    parts.append(np.uint64(0x7800000000000000) | (rng.integers(0, 500, 2000, dtype=np.uint64) << np.uint64(40))
                 | rng.integers(0, 2**40, 2000, dtype=np.uint64))
    # runs of exactly equal keys, longer than the insertion limit, need no fixing
    parts.append(np.uint64(0x7900000000000000) | (rng.integers(0, 20, 1500, dtype=np.uint64) << np.uint64(44)))
    # a run of 9 (just over the limit) in reverse order
    parts.append(np.uint64(0x7A00000000000000) | np.arange(9, 0, -1, dtype=np.uint64))
    # and of 8 (at the limit)
    parts.append(np.uint64(0x7B00000000000000) | np.arange(8, 0, -1, dtype=np.uint64))
    keys = np.concatenate(parts)
    rng.shuffle(keys)
    _check_u64(rt, keys)
    # chained: a second stable sort through perm_in
    perm_in = rng.permutation(len(keys)).astype(np.uint32)
    _check_u64(rt, keys, perm_in=perm_in)


@pytest.mark.parametrize("desc", [False, True])
def test_int64_with_nulls_uses_validity_place(desc):
    rt = dev.default_runtime()
    rng = np.random.default_rng(11)
    n = 700_001
    keys = rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64)
    keys[rng.random(n) < 0.01] = np.iinfo(np.int64).max if not desc else np.iinfo(np.int64).min
    valid = rng.random(n) >= 0.2
    u = keys.view(np.uint64) ^ np.uint64(1 << 63)
    if desc:
        u = ~u
    # nulls after every value: 65-bit key (validity, radix key)
    order = np.lexsort((np.arange(n), np.where(valid, u, 0), ~valid))
    got, _ = _sort(rt, keys, _cabi.I64, desc=desc, validity=valid)
    assert np.array_equal(got, order.astype(np.uint32))
    lsd, _ = _sort(rt, keys, _cabi.I64, desc=desc, validity=valid, algorithm=1)
    assert np.array_equal(lsd, order.astype(np.uint32))


def test_float64_keys_nan_and_signed_zero():
    rt = dev.default_runtime()
    rng = np.random.default_rng(13)
    n = 500_009
    f = rng.standard_normal(n) * 10.0 ** rng.integers(-30, 30, n)
    f[rng.random(n) < 0.02] = np.nan
    f[rng.random(n) < 0.02] = 0.0
    f[rng.random(n) < 0.02] = -0.0
    got, _ = _sort(rt, f, _cabi.F64)
    lsd, _ = _sort(rt, f, _cabi.F64, algorithm=1)
    assert np.array_equal(got, lsd)
    s = f[got]
    finite = s[~np.isnan(s)]
    assert np.all(np.diff(finite) >= 0) and np.all(np.isnan(s[len(finite):]))
    # stability: equal values (and the -0.0 / +0.0 tie) keep row order
    same = s[1:] == s[:-1]
    assert np.all(got[1:][same] > got[:-1][same])

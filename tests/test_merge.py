"""K6, the stable k-way merge (xchg_merge_sorted_runs), through the C-ABI and through
``B200ArrowBlockAccessor.merge_sorted_blocks``.

Reference semantics: ``merge_sorted_blocks`` = concatenate the runs in mapper order and
stable-sort (python/ray/data/_internal/arrow_block.py:529-541 ->
_internal/arrow_ops/transform_pyarrow.py:540-547), so equal keys keep (run index,
position in run) order.  Oracle: oracle.merge_sorted_blocks (that call sequence) and, for
the raw kernel, numpy's stable argsort of the concatenation.
"""
import ctypes

import numpy as np
import pyarrow as pa
import pytest
import torch

from _golden import tables_equal
from cluster_anywhere_b200 import B200ArrowBlockAccessor, SortKey, _cabi
from cluster_anywhere_b200 import device as dev
from oracle import exchange_oracle as O


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _merge(rt, keys, starts, vals=None, want_flag=True):
    n = len(keys)
    k_d = torch.from_numpy(keys.view(np.int64).copy()).to(rt.device) if n else rt.empty(1, torch.int64)
    v_d = torch.from_numpy(vals.view(np.int32).copy()).to(rt.device) if vals is not None else None
    a_st = (ctypes.c_uint64 * len(starts))(*[int(x) for x in starts])
    wsb = ctypes.c_uint64()
    _cabi.check(_cabi.load().xchg_merge_workspace_bytes(n, ctypes.byref(wsb)))
    ws = rt.empty(wsb.value)
    ko = rt.zeros(max(1, n), torch.int64)
    vo = rt.zeros(max(1, n), torch.int32)
    flag = rt.zeros(1, torch.int32)
    rt.call("xchg_merge_sorted_runs", _vp(k_d), _vp(v_d) if v_d is not None else None, a_st, len(starts) - 1,
            _vp(ko), _vp(vo), _vp(ws), wsb.value, _vp(flag) if want_flag else None, rt.stream)
    rt.synchronize()
    return ko.cpu().numpy().view(np.uint64)[:n], vo.cpu().numpy().view(np.uint32)[:n], int(flag.item())


@pytest.mark.parametrize("run_sizes", [
    [0], [1], [5000], [3, 4], [4096, 4096], [4097, 1, 9000], [10000, 0, 0, 7, 12345, 1, 8192, 3],
    [700] * 17, [1] * 33, [20000, 30, 20000, 30, 5], [2500] * 128,
])
@pytest.mark.parametrize("dup", [False, True])
def test_merge_kernel_vs_stable_sort(engine, run_sizes, dup):
    rt = dev.Runtime()
    rng = np.random.default_rng(sum(run_sizes) + len(run_sizes) + int(dup))
    runs = []
    for m in run_sizes:
        k = rng.integers(0, 40, m).astype(np.uint64) if dup else rng.integers(0, 2 ** 64 - 1, m, dtype=np.uint64)
        if dup and m and rng.random() < 0.3:
            k[:] = 7  # a run of one value
        runs.append(np.sort(k))
    keys = np.concatenate(runs) if runs else np.zeros(0, np.uint64)
    starts = np.concatenate([[0], np.cumsum(run_sizes)])
    order = np.argsort(keys, kind="stable")
    ko, vo, flag = _merge(rt, keys, starts)
    assert flag == 0
    assert np.array_equal(ko, keys[order])
    assert np.array_equal(vo, order.astype(np.uint32))  # ties: (run, position) order
    # values travelling with the keys instead of positions
    vals = rng.integers(0, 2 ** 32 - 1, len(keys), dtype=np.uint32)
    ko2, vo2, _ = _merge(rt, keys, starts, vals=vals, want_flag=False)
    assert np.array_equal(ko2, keys[order]) and np.array_equal(vo2, vals[order])


def test_merge_reports_an_unsorted_run(engine):
    rt = dev.Runtime()
    keys = np.array([1, 2, 3, 10, 9, 11, 0, 5], dtype=np.uint64)  # run 1 = [10, 9, 11] is not sorted
    _, _, flag = _merge(rt, keys, [0, 3, 6, 8])
    assert flag == 1
    _, _, flag = _merge(rt, np.array([1, 2, 3, 9, 10, 11, 0, 5], dtype=np.uint64), [0, 3, 6, 8])
    assert flag == 0
    rc = _cabi.load().xchg_merge_sorted_runs(None, None, (ctypes.c_uint64 * 200)(), 199, None, None, None, 0, None,
                                             None)
    assert rc == 0  # 199 empty runs: nothing to do
    big = (ctypes.c_uint64 * 131)(*range(131))
    k = rt.zeros(130, torch.int64)
    v = rt.zeros(130, torch.int32)
    rc = _cabi.load().xchg_merge_sorted_runs(_vp(k), None, big, 130, _vp(k), _vp(v), None, 0, None, rt.stream)
    assert rc == -1 and b"runs per call" in _cabi.load().xchg_last_error()


def _table(rng, n, ktype, nulls):
    if pa.types.is_floating(ktype):
        k = rng.standard_normal(n).round(1)
        k[rng.random(n) < 0.1] = np.nan
    else:
        k = rng.integers(0 if pa.types.is_unsigned_integer(ktype) else -5, 6, n)
    mask = rng.random(n) < 0.15 if nulls else None
    return pa.table({"k": pa.array(k, ktype, mask=mask), "i": pa.array(np.arange(n), pa.int64()),
                     "s": pa.array([None if rng.random() < 0.2 else "x" * int(rng.integers(0, 5)) for _ in range(n)])})


@pytest.mark.parametrize("ktype,nulls", [(pa.int64(), False), (pa.int32(), True), (pa.float64(), True),
                                         (pa.float32(), False), (pa.uint8(), True), (pa.int64(), True)])
@pytest.mark.parametrize("desc", [False, True])
def test_merge_sorted_blocks_vs_oracle(engine, ktype, nulls, desc):
    rng = np.random.default_rng(17)
    sk = SortKey("k", desc)
    runs = []
    for m in (900, 0, 1, 5000, 333):
        t = _table(rng, m, ktype, nulls)
        runs.append(O.sort_table(t, ["k"], [desc]) if m else t)
    got, meta = B200ArrowBlockAccessor.merge_sorted_blocks(runs, sk)
    exp = O.merge_sorted_blocks(runs, ["k"], [desc])
    assert meta.num_rows == exp.num_rows
    assert tables_equal(got, exp)


def test_merge_sorted_blocks_with_an_unsorted_run_still_sorts(engine):
    """The reference re-sorts whatever it is given; a caller that hands over an unsorted
    'run' must get the same answer here (the merge notices and the full sort takes over)."""
    rng = np.random.default_rng(3)
    runs = [_table(rng, 400, pa.int32(), False) for _ in range(3)]  # not sorted at all
    got, _ = B200ArrowBlockAccessor.merge_sorted_blocks(runs, SortKey("k"))
    assert tables_equal(got, O.merge_sorted_blocks(runs, ["k"], [False]))

"""Kernels of the multi-GPU exchange's control path, called through the C-ABI: K1 sampling
(xchg_sample_keys), boundary selection (xchg_select_boundaries), the destination
histogram and stable partition (xchg_partition_hist / xchg_partition_rows), segment copies
and arrival flags.  Expectations are computed with numpy here (independently of the
library and of its host-memory double); on the B200 the sampled ROWS must also equal the
double's, because every rank -- and the CPU tests -- must draw the same samples.

Reference semantics: sample = min(n_samples, rows) DISTINCT rows of every block
(_internal/arrow_block.py:341-344); boundaries = samples at int(q * (len - 1)) for q in
linspace(0, 1, R + 1)[1:-1] (planner/exchange/sort_task_spec.py:221-226); destination =
find_partitions' range partition (_internal/util.py:806-828).
"""
import ctypes

import numpy as np
import pytest
import torch

from _fake_cabi import FakeLib, _normalize
from cluster_anywhere_b200 import _cabi
from cluster_anywhere_b200 import device as dev

_NP = {_cabi.I8: np.int8, _cabi.I16: np.int16, _cabi.I32: np.int32, _cabi.I64: np.int64, _cabi.U8: np.uint8,
       _cabi.U16: np.uint16, _cabi.U32: np.uint32, _cabi.U64: np.uint64, _cabi.F32: np.float32,
       _cabi.F64: np.float64}


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _to_dev(rt, a):
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy())
    return t.to(rt.device)


def _keys(rng, kt, n):
    dt = _NP[kt]
    if np.dtype(dt).kind == "f":
        k = (rng.standard_normal(n) * 100).astype(dt)
        k[rng.random(n) < 0.05] = np.nan
        k[rng.random(n) < 0.05] = 0.0
        k[rng.random(n) < 0.05] = -0.0
        return k
    info = np.iinfo(dt)
    if rng.random() < 0.5:
        return rng.integers(max(info.min, -40), min(info.max, 40), n, endpoint=True).astype(dt)  # duplicates
    return rng.integers(info.min, info.max, n, endpoint=True, dtype=dt)


@pytest.mark.parametrize("kt", [_cabi.I64, _cabi.I32, _cabi.U8, _cabi.F32, _cabi.F64])
def test_sample_keys(engine, kt):
    rt = dev.Runtime()
    rng = np.random.default_rng(kt)
    rows = [1000, 0, 3, 10, 25000, 7]
    starts = np.concatenate([[0], np.cumsum(rows)])[:-1]
    n = int(sum(rows))
    keys = _keys(rng, kt, n)
    keys_d = _to_dev(rt, keys)
    for desc in (0, 1):
        cap = 64
        out = rt.zeros(cap, torch.int64)
        cnt = ctypes.c_uint64()
        a_st = (ctypes.c_uint64 * len(rows))(*[int(x) for x in starts])
        a_rw = (ctypes.c_uint64 * len(rows))(*rows)
        rt.call("xchg_sample_keys", _vp(keys_d), kt, desc, None, 0, a_st, a_rw, len(rows), 10, 77, 5, _vp(out),
                cap, ctypes.byref(cnt), rt.stream)
        rt.synchronize()
        exp_cnt = sum(min(10, r) for r in rows)
        assert cnt.value == exp_cnt
        got = out.cpu().numpy().view(np.uint64)[:exp_cnt]
        nk = _normalize(keys, kt, bool(desc))
        pos = 0
        for b, r in enumerate(rows):
            k = min(10, r)
            seg = got[pos:pos + k]
            block_keys = nk[starts[b]:starts[b] + r]
            # every sample is a key of ITS block, and k distinct ROWS were drawn: as a
            # multiset the samples fit into the block's keys
            avail = {}
            for x in block_keys.tolist():
                avail[x] = avail.get(x, 0) + 1
            for x in seg.tolist():
                assert avail.get(x, 0) > 0
                avail[x] -= 1
            if k == r:
                assert sorted(seg.tolist()) == sorted(block_keys.tolist())
            pos += k
        # the double draws the same rows (host logic and tests see the device's samples)
        if engine == "b200":
            f_out = np.zeros(cap, dtype=np.uint64)
            f_cnt = ctypes.c_uint64()
            FakeLib().xchg_sample_keys(ctypes.c_void_p(keys.ctypes.data), kt, desc, None, 0, a_st, a_rw, len(rows), 10,
                                       77, 5, ctypes.c_void_p(f_out.ctypes.data), cap, ctypes.byref(f_cnt), None)
            assert f_cnt.value == exp_cnt and np.array_equal(f_out[:exp_cnt], got)
        # capacity is enforced
        rc = _cabi.load().xchg_sample_keys(_vp(keys_d), kt, desc, None, 0, a_st, a_rw, len(rows), 10, 77, 5, _vp(out),
                                           exp_cnt - 1, ctypes.byref(cnt), rt.stream)
        assert rc == -4


@pytest.mark.parametrize("counts,R", [([7, 0, 120, 33], 12), ([1], 5), ([0, 0], 3), ([4096, 4096, 4096, 4096, 4096], 96),
                                      ([500, 500], 1)])
def test_select_boundaries(engine, counts, R):
    rt = dev.Runtime()
    rng = np.random.default_rng(len(counts) * 100 + R)
    HDR, CAP = 8, 4096
    msgs = np.zeros((len(counts), HDR + CAP), dtype=np.uint64)
    for g, c in enumerate(counts):
        msgs[g, 3] = c
        msgs[g, HDR:HDR + c] = rng.integers(0, 2 ** 64 - 1, c, dtype=np.uint64) if rng.random() < 0.7 else \
            rng.integers(0, 5, c).astype(np.uint64)
        msgs[g, HDR + c:] = 12345  # stale slots must be ignored
    g_d = _to_dev(rt, msgs)
    bounds = rt.zeros(max(1, R - 1), torch.int64)
    info = rt.zeros(2, torch.int64)
    rt.call("xchg_select_boundaries", _vp(g_d), len(counts), HDR + CAP, 3, HDR, R, _vp(bounds), _vp(info), rt.stream)
    rt.synchronize()
    allv = np.concatenate([msgs[g, HDR:HDR + c] for g, c in enumerate(counts)])
    total = len(allv)
    stride = -(-total // 16384) if total > 16384 else 1
    used = np.sort(allv[::stride])
    assert info.cpu().numpy().tolist() == [total, len(used)]
    if len(used) and R > 1:
        qs = np.linspace(0, 1, R + 1)[1:-1]
        exp = np.array([used[int(q * (len(used) - 1))] for q in qs], dtype=np.uint64)
        assert np.array_equal(bounds.cpu().numpy().view(np.uint64)[:R - 1], exp)


def _expected_dest(nk, thr, strict):
    d = np.zeros(len(nk), dtype=np.int64)
    for t, s in zip(thr.tolist(), strict):
        d += (nk > t) | ((nk == t) & (not s))
    return d


@pytest.mark.parametrize("kt", [_cabi.I64, _cabi.I32, _cabi.I16, _cabi.I8, _cabi.U64, _cabi.U32, _cabi.U16, _cabi.U8,
                                _cabi.F32, _cabi.F64])
@pytest.mark.parametrize("n,nthr", [(0, 3), (1, 0), (5000, 1), (4096 * 3, 7), (70001, 31), (20000, 255)])
def test_partition_hist_and_rows(engine, kt, n, nthr):
    rt = dev.Runtime()
    rng = np.random.default_rng(kt * 1000 + n + nthr)
    keys = _keys(rng, kt, n)
    keys_d = _to_dev(rt, keys) if n else rt.empty(8)
    for desc in (0, 1):
        nk = _normalize(keys, kt, bool(desc)) if n else np.zeros(0, np.uint64)
        # thresholds: key values that occur, with repeats, in sort order; scattered in a
        # longer boundary list and addressed through `pick`
        pool = np.sort(nk) if n else np.arange(10, dtype=np.uint64)
        thr = np.sort(pool[rng.integers(0, len(pool), nthr)]) if nthr else np.zeros(0, np.uint64)
        strict = [int(desc) if rng.random() < 0.8 else 1 - int(desc) for _ in range(nthr)]
        # the rule must be a prefix: among equal thresholds the non-strict ones come first
        order = sorted(range(nthr), key=lambda i: (int(thr[i]), strict[i]))
        thr = thr[order] if nthr else thr
        strict = [strict[i] for i in order]
        bounds = np.zeros(2 * nthr + 1, dtype=np.uint64)
        pick = [2 * i for i in range(nthr)]
        bounds[pick] = thr
        b_d = _to_dev(rt, bounds)
        a_pick = (ctypes.c_uint32 * max(1, nthr))(*pick)
        a_strict = (ctypes.c_uint8 * max(1, nthr))(*strict)
        counts = rt.zeros(nthr + 1, torch.int64)
        rt.call("xchg_partition_hist", _vp(keys_d), kt, desc, n, _vp(b_d), a_pick, a_strict, nthr, _vp(counts),
                rt.stream)
        exp_d = _expected_dest(nk, thr, strict)
        exp_counts = np.bincount(exp_d, minlength=nthr + 1)
        rt.synchronize()
        assert np.array_equal(counts.cpu().numpy(), exp_counts)
        wsb = ctypes.c_uint64()
        _cabi.check(_cabi.load().xchg_partition_workspace_bytes(n, ctypes.byref(wsb)))
        ws = rt.empty(wsb.value)
        ko = rt.zeros(max(1, n), torch.int64)
        ro = rt.zeros(max(1, n), torch.int32)
        rt.call("xchg_partition_rows", _vp(keys_d), kt, desc, n, _vp(b_d), a_pick, a_strict, nthr, _vp(counts), _vp(ko),
                _vp(ro), _vp(ws), wsb.value, rt.stream)
        rt.synchronize()
        order = np.argsort(exp_d, kind="stable")
        assert np.array_equal(ro.cpu().numpy().view(np.uint32)[:n], order.astype(np.uint32))
        assert np.array_equal(ko.cpu().numpy().view(np.uint64)[:n], nk[order])


def test_copy_segments(engine):
    rt = dev.Runtime()
    rng = np.random.default_rng(4)
    src = _to_dev(rt, rng.integers(0, 256, 1 << 20, dtype=np.uint8))
    dst = rt.zeros(1 << 21)
    segs = [(0, 0, 100000), (100003, 200001, 65536 + 5), (300000, 400008, 8 * 777), (500016, 600032, 32768 * 3 + 16),
            (0, 900000, 0), (700001, 1000000, 1)] + [(1000 * i, 1100000 + 3000 * i, 999) for i in range(20)]
    a_src = (ctypes.c_void_p * len(segs))(*[src.data_ptr() + s for s, d, b in segs])
    a_dst = (ctypes.c_void_p * len(segs))(*[dst.data_ptr() + d for s, d, b in segs])
    a_nb = (ctypes.c_uint64 * len(segs))(*[b for s, d, b in segs])
    rt.call("xchg_copy_segments", len(segs), a_src, a_dst, a_nb, rt.stream)
    rt.synchronize()
    s_h, d_h = src.cpu().numpy(), dst.cpu().numpy()
    exp = np.zeros(1 << 21, dtype=np.uint8)
    for s, d, b in segs:
        exp[d:d + b] = s_h[s:s + b]
    assert np.array_equal(d_h, exp)


def test_arrival_flags(engine):
    rt = dev.Runtime()
    flags = rt.zeros(16, torch.int32)
    err = rt.zeros(1, torch.int32)
    ptrs = (ctypes.c_void_p * 3)(*[flags.data_ptr() + 4 * i for i in (2, 3, 4)])
    rt.call("xchg_signal_flags", ptrs, 3, 5, rt.stream)
    rt.call("xchg_wait_flags", ctypes.c_void_p(flags.data_ptr() + 8), 3, 5, 10 ** 9, _vp(err), rt.stream)
    rt.call("xchg_wait_flags", ctypes.c_void_p(flags.data_ptr() + 8), 3, 4, 10 ** 9, _vp(err), rt.stream)  # already past
    rt.synchronize()
    assert flags.cpu().numpy().tolist()[:6] == [0, 0, 5, 5, 5, 0] and int(err.item()) == 0
    # a flag that never arrives: the wait gives up after its timeout and says which flag
    rt.call("xchg_wait_flags", ctypes.c_void_p(flags.data_ptr() + 8), 4, 5, 20 * 10 ** 6, _vp(err), rt.stream)
    rt.synchronize()
    assert int(err.item()) == 4  # 1 + index of the missing flag

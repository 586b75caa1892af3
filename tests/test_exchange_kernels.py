"""Exchange-side kernels: per-row range partitioning (xchg_assign_partitions), bitmap
<-> byte conversion, offsets re-basing, and the whole map -> all-to-all -> reduce path
on ONE rank (a world-size-1 process group: NCCL on the GPU, gloo on the host double).

Oracle for the per-row rule: sort with the oracle, cut with its find_partitions
(the reference's own rule, util.py:806-828), read every row's partition back
through a row-id column."""
import os
import socket

import numpy as np
import pyarrow as pa
import pytest
import torch

import _golden as G
from cluster_anywhere_b200 import SortKey
from cluster_anywhere_b200 import device as dev
from oracle import exchange_oracle as O


def _table(rng, n):
    k = rng.integers(-8, 8, n)
    f = np.round(rng.standard_normal(n) * 2) / 2
    f[rng.random(n) < 0.05] = np.nan
    return pa.table({
        "k": pa.array(k, pa.int64(), mask=rng.random(n) < 0.1),
        "f": pa.array(f, pa.float32(), mask=rng.random(n) < 0.1),
        "u": pa.array(rng.integers(0, 5, n), pa.uint8()),
        "g": pa.array(np.round(rng.standard_normal(n) * 2) / 2, pa.float64()),
        "row": pa.array(np.arange(n, dtype=np.int64)),
    })


_CASES = [
    (["k"], [False], [(-3,), (0,), (0,), (4,)]),
    (["k"], [True], [(5,), (0,), (-2.5,)]),
    (["f"], [False], [(-0.5,), (0.25,), (1.0,)]),
    (["u"], [False], [(1,), (2,), (7,)]),
    # a DESCENDING column with NaN / null rows is where the reference's cut points follow
    # numpy's probe sequence over an unsorted view (csrc/partition.cu): no per-row rule
    # exists there, so descending secondary keys are NaN- and null-free here
    (["k", "g"], [False, True], [(-3, 1.0), (0, None), (0, 0.5), (0, -0.5), (4, 0.0)]),
    (["k", "f"], [False, False], [(-3, 1.0), (0, -0.5), (0, 0.5), (0, None), (4, 0.0)]),
    (["k", "u"], [True, False], [(5, 1), (0, 3), (0,), (-4, 2)]),
    (["u", "k", "f"], [False, False, False], [(1, 0, 0.0), (1, 0, 0.5), (3, -2, -1.0)]),
    (["k"], [False], [(None,)]),
    (["k"], [False], []),
]


@pytest.mark.parametrize("case", range(len(_CASES)))
def test_assign_partitions_matches_find_partitions(engine, case):
    cols, desc, bounds = _CASES[case]
    rng = np.random.default_rng(50 + case)
    n = 30011 if engine == "b200" else 2000
    t = _table(rng, n)
    rt = dev.default_runtime()
    dt = dev.stage_table(rt, t)
    sk = SortKey(cols, desc)
    dest, counts = dev.assign_partitions(rt, dt, bounds, sk)
    rt.synchronize()
    got = dest.cpu().numpy().view(np.uint32)
    parts = O.sort_and_partition(t, bounds, cols, desc)
    exp = np.empty(n, dtype=np.uint32)
    for p, part in enumerate(parts):
        if part.num_rows:
            exp[part["row"].to_numpy()] = p
    assert np.array_equal(got, exp)
    assert counts.cpu().numpy().tolist() == [p.num_rows for p in parts]
    # the stable order by destination keeps the original row order inside a partition
    perm = dev.sort_u32(rt, dest, n).cpu().numpy().view(np.uint32)
    assert np.array_equal(perm, np.argsort(exp, kind="stable").astype(np.uint32))


def test_bitmap_and_offsets_utilities(engine):
    rt = dev.default_runtime()
    rng = np.random.default_rng(5)
    for n, bit0 in ((1, 0), (31, 3), (32, 0), (33, 5), (100_003, 7)):
        bits = rng.random(n + bit0) < 0.3
        packed = np.packbits(bits, bitorder="little")
        src = rt.zeros(len(packed) + 8)
        dev._h2d(src[: len(packed)], packed)
        by = rt.empty(n)
        rt.call("xchg_bitmap_unpack", dev._vp(src), bit0, n, dev._vp(by), rt.stream)
        out = rt.zeros(((n + 31) // 32) * 4)
        zeros = rt.zeros(1, torch.int64)
        rt.call("xchg_bitmap_pack", dev._vp(by), n, dev._vp(out), dev._vp(zeros), rt.stream)
        rt.synchronize()
        want = bits[bit0:bit0 + n]
        assert np.array_equal(by.cpu().numpy().astype(bool), want)
        assert np.array_equal(np.unpackbits(out.cpu().numpy(), bitorder="little")[:n].astype(bool), want)
        assert int(zeros.item()) == int((~want).sum())
    off = np.concatenate([[1000], 1000 + np.cumsum(rng.integers(0, 50, 70_001))]).astype(np.int64)
    src = rt.empty(len(off), torch.int64)
    dev._h2d(src.view(torch.uint8), off.view(np.uint8))
    dst = rt.empty(len(off), torch.int64)
    rt.call("xchg_offsets_rebase", dev._vp(src), len(off), 77, dev._vp(dst), rt.stream)
    rt.synchronize()
    assert np.array_equal(dst.cpu().numpy(), off - 1000 + 77)


def _single_rank_group(backend):
    import torch.distributed as dist

    if dist.is_initialized():
        return
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ["MASTER_PORT"] = str(port)
    kw = {}
    if backend == "nccl":
        kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group(backend, rank=0, world_size=1, **kw)


@pytest.mark.parametrize("cols,desc", [(["k"], [False]), (["k", "f"], [True, False]), (["f"], [True])])
def test_single_rank_map_exchange_reduce(engine, cols, desc):
    """force_exchange=True drives assign -> order by destination -> gather -> all-to-all-v
    -> receive-side sort + cuts even with one rank: same kernels and collectives as N>1."""
    import torch.distributed as dist

    from cluster_anywhere_b200.exchange import DeviceExchange

    _single_rank_group("nccl" if engine == "b200" else "gloo")
    try:
        rng = np.random.default_rng(11)
        sizes = (5000, 0, 1203, 9000) if engine == "b200" else (700, 0, 300, 900)
        blocks = []
        for n in sizes:
            t = _table(rng, n)
            s = [None if rng.random() < 0.1 else "v" * int(rng.integers(0, 20)) for _ in range(n)]
            t = t.append_column("s", pa.array(s, pa.string()))
            t = t.append_column("b", pa.array(rng.random(n) < 0.5))
            t = t.append_column("p", pa.array([bytes(rng.integers(0, 256, 12, dtype=np.uint8))
                                               for _ in range(n)], pa.binary()))
            # 32-byte rows: the 16-byte-vector variant of the (multi-destination) row gather
            t = t.append_column("q", pa.array([bytes(rng.integers(0, 256, 32, dtype=np.uint8))
                                               for _ in range(n)], pa.binary()))
            blocks.append(t)
        ex = DeviceExchange()
        sk = SortKey(cols, desc)
        out = ex.export(ex.sort_distributed(ex.stage(blocks), sk, seed=5, force_exchange=True))
        assert len(out) == len(blocks)
        exp = O.sort_table(pa.concat_tables([b for b in blocks if b.num_rows]), cols, desc)
        assert G.tables_equal(pa.concat_tables([o for o in out if o.num_rows]), exp)
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("w,n_in,n_out", [(1024, 5000, 7001), (16, 100, 3), (4096, 300, 299), (48, 70000, 100000)])
def test_gather_rows_tma_variant_equals_the_default(w, n_in, n_out):
    """XCHG_GATHER_BULK=1 sends 16-byte-aligned uniform rows through cp.async.bulk (global ->
    shared -> global, no thread touches the data); same bytes as the default kernel."""
    import ctypes

    from cluster_anywhere_b200 import _cabi

    _cabi._install_test_double(None)
    dev.reset_default_runtime()
    rt = dev.default_runtime()
    rng = np.random.default_rng(w)
    src = torch.from_numpy(rng.integers(0, 256, n_in * w, dtype=np.uint8)).to(rt.device)
    idx = torch.from_numpy(rng.integers(0, n_in, n_out).astype(np.int32)).to(rt.device)
    libc = ctypes.CDLL(None)
    outs = []
    for mode in (b"0", b"1"):
        libc.setenv(b"XCHG_GATHER_BULK", mode, 1)
        out = rt.zeros(n_out * w)
        rt.call("xchg_gather_rows", ctypes.c_void_p(src.data_ptr()), w, ctypes.c_void_p(idx.data_ptr()), n_out,
                ctypes.c_void_p(out.data_ptr()), rt.stream)
        rt.synchronize()
        outs.append(out.cpu().numpy())
    libc.setenv(b"XCHG_GATHER_BULK", b"0", 1)
    exp = src.cpu().numpy().reshape(n_in, w)[idx.cpu().numpy()].reshape(-1)
    assert np.array_equal(outs[0], exp) and np.array_equal(outs[1], exp)


@pytest.mark.parametrize("kind", ["short", "mixed", "huge_row", "empties"])
def test_ragged_take_with_cached_spans_equals_the_plain_take(engine, kind):
    """xchg_gather_spans + xchg_gather_bytes_spans (source starts kept between the offsets
    pass and the byte copy) against xchg_gather_offsets + xchg_gather_bytes, and both against
    numpy: identical offsets and bytes (Table.take on a var-width column,
    transform_pyarrow.py:45-71)."""
    import ctypes

    from cluster_anywhere_b200 import _cabi

    rt = dev.Runtime()
    rng = np.random.default_rng({"short": 1, "mixed": 2, "huge_row": 3, "empties": 4}[kind])
    n_in = 3000 if engine == "b200" else 300
    if kind == "short":
        lens = rng.integers(0, 24, n_in)
    elif kind == "mixed":
        lens = rng.integers(0, 700, n_in)
    elif kind == "huge_row":
        lens = rng.integers(0, 90, n_in)
        lens[rng.integers(0, n_in, 3)] = 70_000  # wider than the staging tile: the whole-CTA copy
    else:
        lens = np.where(rng.random(n_in) < 0.8, 0, rng.integers(1, 40, n_in))
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    data = rng.integers(0, 256, int(off[-1]) + 16, dtype=np.uint8)
    n_out = 2 * n_in + 37
    idx = rng.integers(0, n_in, n_out).astype(np.uint32)

    def up(a, dt):
        t = rt.empty(max(1, a.size), dt)
        if a.size:
            dev._h2d(t.view(torch.uint8)[: a.nbytes], np.ascontiguousarray(a).view(np.uint8).reshape(-1))
        return t

    d_off, d_data, d_idx = up(off, torch.int64), up(data, torch.uint8), up(idx.view(np.int32), torch.int32)
    wsb = ctypes.c_uint64()
    _cabi.check(_cabi.load().xchg_gather_offsets_workspace_bytes(n_out, ctypes.byref(wsb)))
    ws = rt.empty(wsb.value)
    vp = dev._vp
    oo1, oo2 = rt.empty(n_out + 1, torch.int64), rt.empty(n_out + 1, torch.int64)
    starts = rt.empty(n_out, torch.int64)
    rt.call("xchg_gather_offsets", vp(d_off), vp(d_idx), n_out, vp(oo1), vp(ws), wsb.value, rt.stream)
    rt.call("xchg_gather_spans", vp(d_off), vp(d_idx), n_out, vp(starts), vp(oo2), vp(ws), wsb.value, rt.stream)
    rt.synchronize()
    exp_off = np.concatenate([[0], np.cumsum(lens[idx])]).astype(np.int64)
    assert np.array_equal(oo1.cpu().numpy(), exp_off) and np.array_equal(oo2.cpu().numpy(), exp_off)
    assert np.array_equal(starts.cpu().numpy(), off[idx])
    total = int(exp_off[-1])
    out1, out2 = rt.zeros(total + 16), rt.zeros(total + 16)
    rt.call("xchg_gather_bytes", vp(d_data), vp(d_off), vp(d_idx), n_out, vp(oo1), vp(out1), rt.stream)
    rt.call("xchg_gather_bytes_spans", vp(d_data), vp(starts), n_out, vp(oo2), vp(out2), rt.stream)
    rt.synchronize()
    exp = np.concatenate([data[off[r]:off[r + 1]] for r in idx]) if total else np.zeros(0, np.uint8)
    assert np.array_equal(out1.cpu().numpy()[:total], exp)
    assert np.array_equal(out2.cpu().numpy()[:total], exp)
    assert not out2.cpu().numpy()[total:].any()  # nothing written past the end

"""The multi-GPU sort's chunked fast path (exchange/chunked_sort.py) on G VIRTUAL ranks:
G threads of one process, each with its own streams and its own arena, all on one device
(exchange/comm.py).  The kernels, arena layouts with non-zero source prefixes, chunk-wise
stores and arrival flags are the production ones; only the three small collectives are
replaced by in-process hand-offs.  Runs on the host-memory double (CPU suite) and on the
real library on a single B200 (-m gpu), which is where the driver can see N > 1 parity.

Oracle: oracle.sort_table / oracle.find_partitions (the reference's call sequence);
reference semantics pinned here: global stable order with ties in (block, position)
order, block j = the j-th range partition under the boundaries.
"""
import numpy as np
import pyarrow as pa
import pytest

from _golden import tables_equal
from cluster_anywhere_b200 import SortKey
from cluster_anywhere_b200 import device as dev
from cluster_anywhere_b200.exchange import DeviceExchange
from cluster_anywhere_b200.exchange.comm import run_virtual_world
from oracle import exchange_oracle as O


def _device(engine):
    if engine == "b200":
        import torch

        return torch.device("cuda", torch.cuda.current_device())
    return None


def _block(rng, n, ktype, lo, hi, width=10):
    if pa.types.is_floating(ktype):
        k = rng.standard_normal(n) * (hi - lo)
        k[rng.random(n) < 0.05] = np.nan
        k[rng.random(n) < 0.05] = 0.0
        k[rng.random(n) < 0.03] = -0.0
    else:
        k = rng.integers(lo, hi, n)
    pay = rng.integers(0, 256, n * width, dtype=np.uint8)
    offs = np.arange(n + 1, dtype=np.int32) * width
    p = pa.Array.from_buffers(pa.binary(), n, [None, pa.py_buffer(offs), pa.py_buffer(pay)])
    return pa.table({"k": pa.array(k, ktype), "v": pa.array(rng.integers(0, 1 << 30, n), pa.int32()), "p": p})


def _run(engine, blocks, sk, boundaries=None, seed=3, chunks=None, monkeypatch=None):
    G = len(blocks)
    if chunks is not None:
        monkeypatch.setenv("XCHG_CHUNKS", str(chunks))

    def fn(comm, r):
        ex = DeviceExchange(dev.Runtime())
        out = ex.sort_blocks_distributed(blocks[r], sk, boundaries, seed, group=comm)
        return out, getattr(ex, "chunked_steps", 0)

    res = run_virtual_world(G, fn, _device(engine))
    return [b for out, _ in res for b in out], [c for _, c in res]


CASES = [
    # (key type, lo, hi, descending, block sizes per rank)
    (pa.int64(), -50, 50, False, [[400, 30, 300], [1, 900], [17, 5, 5, 250]]),
    (pa.int64(), -50, 50, True, [[400, 30, 300], [1, 900], [17, 5, 5, 250]]),
    (pa.int32(), -(2 ** 31), 2 ** 31 - 1, False, [[5000, 7000], [6000, 1]]),
    (pa.uint8(), 0, 256, True, [[3000], [10], [2000, 2000], [50, 60, 70]]),
    (pa.uint64(), 0, 2 ** 63, False, [[9000, 100], [100, 9000]]),
    (pa.int16(), -7, 8, False, [[2500, 2500, 2500, 2500, 2500], [1, 1, 1, 1, 1]]),
    (pa.float64(), -3, 3, False, [[3000, 500], [2500, 700, 9]]),
    (pa.float32(), -3, 3, False, [[1500, 1500], [1500, 1500]]),
]


@pytest.mark.parametrize("ktype,lo,hi,desc,sizes", CASES)
def test_sampled_boundaries_global_order(engine, ktype, lo, hi, desc, sizes, monkeypatch):
    rng = np.random.default_rng(11)
    blocks = [[_block(rng, n, ktype, lo, hi) for n in s] for s in sizes]
    full = pa.concat_tables([b for bs in blocks for b in bs])
    sk = SortKey("k", desc)
    for chunks in (1, 3):
        got, used = _run(engine, blocks, sk, chunks=chunks, monkeypatch=monkeypatch)
        assert used == [1] * len(blocks), "the chunked path must have run on every rank"
        assert len(got) == sum(len(s) for s in sizes)  # R = number of input blocks
        exp = O.sort_table(full, ["k"], [desc])
        assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), exp)
        # every block is a key range: no key of block j sorts after a key of block j + 1
        nonempty = [g for g in got if g.num_rows]
        cat = pa.concat_tables(nonempty)
        assert cat.num_rows == full.num_rows


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("bounds", [[(-20,), (0,), (0,), (31,)], [(-1000,), (7,)], [(2.5,), (10.0,), (1e9,)]])
def test_user_boundaries_block_contents(engine, desc, bounds, monkeypatch):
    rng = np.random.default_rng(5)
    sizes = [[800, 10], [700, 300, 2]]
    blocks = [[_block(rng, n, pa.int32(), -50, 50) for n in s] for s in sizes]
    full = pa.concat_tables([b for bs in blocks for b in bs])
    sk = SortKey("k", desc)
    bnds = sorted(bounds, reverse=desc)  # planner/sort.py:59-62 hands them over in sort order
    got, used = _run(engine, blocks, sk, boundaries=bnds, chunks=2, monkeypatch=monkeypatch)
    assert used == [1, 1]
    exp = O.sort_table(full, ["k"], [desc])
    parts = O.find_partitions(exp, bnds, ["k"], [desc])
    assert len(got) == len(parts)
    for g, e in zip(got, parts):
        assert g.num_rows == e.num_rows
        if e.num_rows:
            assert tables_equal(g, e)


def test_all_equal_keys_land_in_one_block(engine, monkeypatch):
    """D/tests/test_sort.py:99-114: equal keys are never split."""
    n = 4000
    blocks = [[pa.table({"k": pa.array(np.full(n, 7), pa.int64()), "v": pa.array(np.arange(n) + r * n, pa.int32())})
               for _ in range(2)] for r in range(2)]
    got, used = _run(engine, blocks, SortKey("k"), chunks=2, monkeypatch=monkeypatch)
    assert used == [1, 1]
    rows = [g.num_rows for g in got]
    assert sorted(rows) == [0, 0, 0, 4 * n]
    full = pa.concat_tables([b for bs in blocks for b in bs])
    assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), full)  # stable: input order


def test_ineligible_input_takes_the_general_path_on_every_rank(engine, monkeypatch):
    rng = np.random.default_rng(2)
    # rank 1's key column has nulls: no rank may take the chunked path
    b0 = _block(rng, 500, pa.int64(), -9, 9)
    b1 = _block(rng, 700, pa.int64(), -9, 9)
    k = b1["k"].to_numpy().copy()
    b1 = b1.set_column(0, "k", pa.array(k, pa.int64(), mask=rng.random(700) < 0.1))
    b0 = b0.set_column(0, "k", pa.array(b0["k"].to_numpy(), pa.int64(), mask=np.zeros(500, bool)))
    blocks = [[b0], [b1]]
    got, used = _run(engine, blocks, SortKey("k"), chunks=2, monkeypatch=monkeypatch)
    assert used == [0, 0]
    exp = O.sort_table(pa.concat_tables([b0, b1]), ["k"], [False])
    assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), exp)


def test_repeated_steps_reuse_the_arena(engine, monkeypatch):
    """Arrival flags carry the step number: a second and third exchange through the same
    arena must not see the first one's flags as theirs."""
    monkeypatch.setenv("XCHG_CHUNKS", "4")
    rng = np.random.default_rng(9)
    sizes = [[1200, 800, 100, 50], [900, 1100, 5, 5]]
    blocks = [[_block(rng, n, pa.int64(), -10 ** 6, 10 ** 6) for n in s] for s in sizes]
    full = pa.concat_tables([b for bs in blocks for b in bs])
    sk = SortKey("k")

    def fn(comm, r):
        ex = DeviceExchange(dev.Runtime())
        outs = [ex.sort_blocks_distributed(blocks[r], sk, None, 100 + i, group=comm) for i in range(3)]
        return outs, ex.chunked_steps

    res = run_virtual_world(2, fn, _device(engine))
    assert [c for _, c in res] == [3, 3]
    exp = O.sort_table(full, ["k"], [False])
    for i in range(3):
        got = [b for outs, _ in res for b in outs[i]]
        assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), exp)

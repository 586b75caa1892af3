"""Device-resident Dataset.sort (DeviceExchange) vs the oracle's pull-based
map/reduce schedule: identical output blocks for identical boundaries, and the
size-independent properties at larger sizes."""
import numpy as np
import pyarrow as pa
import pytest

import _golden as G
from cluster_anywhere_b200 import SortKey
from cluster_anywhere_b200.exchange import DeviceExchange
from oracle import exchange_oracle as O


@pytest.mark.parametrize("case", G.cases("sort"), ids=lambda c: c["name"])
def test_golden_sort_device_exchange(engine, case):
    blocks = [G.load(n) for n in case["inputs"]]
    b = G.boundaries(case)
    sk = SortKey(case["columns"], case["descending"])
    outs = DeviceExchange().sort_blocks(blocks, sk, boundaries=b)
    assert len(outs) == len(case["outputs"])
    for o, gold in zip(outs, case["outputs"]):
        assert G.tables_equal(o, G.load(gold)), gold


def _gray_blocks(rng, sizes, dup_keys=False):
    blocks = []
    for n in sizes:
        hi = 50 if dup_keys else 2**63 - 1
        keys = rng.integers(0, hi, n, dtype=np.int64)
        payload = rng.integers(0, 256, n * 90, dtype=np.uint8)
        offs = np.arange(n + 1, dtype=np.int32) * 90
        c1 = pa.Array.from_buffers(pa.binary(), n, [None, pa.py_buffer(offs), pa.py_buffer(payload)])
        blocks.append(pa.table({"c_0": keys, "c_1": c1}))
    return blocks


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("dup", [False, True])
def test_sampled_boundaries_sort(engine, desc, dup):
    rng = np.random.default_rng(7)
    sizes = [30000, 0, 41000, 29000] if engine == "b200" else [900, 0, 700, 500]
    blocks = _gray_blocks(rng, sizes, dup)
    sk = SortKey("c_0", desc)
    ex = DeviceExchange()
    ds = ex.stage(blocks)
    bounds = ex.sample_boundaries(ds, sk, len(blocks), seed=3)
    assert len(bounds) == len(blocks) - 1
    assert bounds == ex.sample_boundaries(ds, sk, len(blocks), seed=3)  # seeded = reproducible
    outs = ex.export(ex.sort(ds, sk, boundaries=bounds))
    exp = O.sort_exchange(blocks, bounds, ["c_0"], [desc])
    assert len(outs) == len(exp) == len(blocks)
    for o, e in zip(outs, exp):
        assert G.tables_equal(o, e)
    # default path: R = number of input blocks, boundaries sampled internally
    outs2 = ex.sort_blocks(blocks, sk, seed=3)
    for o, e in zip(outs2, exp):
        assert G.tables_equal(o, e)


def test_all_equal_keys_land_in_one_block(engine):
    # python/ray/data/tests/test_sort.py:99-114
    blocks = [pa.table({"k": np.full(500, 7, dtype=np.int64), "i": np.arange(500) + 500 * b})
              for b in range(4)]
    outs = DeviceExchange().sort_blocks(blocks, SortKey("k"))
    sizes = [o.num_rows for o in outs]
    assert sorted(sizes) == [0, 0, 0, 2000]
    full = [o for o in outs if o.num_rows][0]
    assert full["i"].to_pylist() == list(range(2000))  # stable across blocks


def test_empty_dataset(engine):
    e = pa.table({"k": pa.array([], pa.int64())})
    outs = DeviceExchange().sort_blocks([e, e, e], SortKey("k"))
    assert len(outs) == 3 and all(o.num_rows == 0 for o in outs)


@pytest.mark.gpu
def test_large_sort_properties():
    """Full-size-style check through size-independent properties (8M GraySort rows:
    ~980 tiles per pass, several waves of CTAs, 8 radix passes): sortedness,
    stability on duplicate keys, row multiset (key + payload checksum pairs)."""
    import torch

    from cluster_anywhere_b200 import _cabi, device

    _cabi._install_test_double(None)
    device.reset_default_runtime()
    n = 8_000_000
    rng = np.random.default_rng(11)
    keys = rng.integers(0, 2**63 - 1, n, dtype=np.int64)
    keys[rng.integers(0, n, n // 10)] = 12345  # a heavy duplicate
    payload = np.empty((n, 90), dtype=np.uint8)
    payload[:] = (np.arange(n, dtype=np.uint64)[:, None] * 2654435761 >> np.arange(90, dtype=np.uint64) % 24).astype(np.uint8)
    payload[:, :8] = np.arange(n, dtype=np.int64).view(np.uint8).reshape(n, 8)  # row id inside the payload
    offs = np.arange(n + 1, dtype=np.int32) * 90
    c1 = pa.Array.from_buffers(pa.binary(), n, [None, pa.py_buffer(offs), pa.py_buffer(payload.reshape(-1))])
    m = 8
    t = pa.table({"c_0": keys, "c_1": c1})
    blocks = [t.slice(i * n // m, n // m) for i in range(m)]
    outs = DeviceExchange().sort_blocks(blocks, SortKey("c_0"))
    res = pa.concat_tables(outs).combine_chunks()
    assert res.num_rows == n
    k = res["c_0"].to_numpy()
    assert (np.diff(k) >= 0).all()
    data = np.frombuffer(res["c_1"].chunk(0).buffers()[2], dtype=np.uint8)[: n * 90].reshape(n, 90)
    rid = data[:, :8].copy().view(np.int64).reshape(n)
    assert (np.sort(rid) == np.arange(n)).all()  # a permutation: multiset preserved
    assert (keys[rid] == k).all()  # every key travelled with its own payload
    assert (data == payload[rid]).all()
    same = np.diff(k) == 0
    assert (np.diff(rid)[same] > 0).all()  # ties keep (block, row) order


def test_export_chunks_a_block_beyond_int32_offsets(engine, monkeypatch):
    """A uniform-width binary block whose bytes exceed the int32 offset range leaves as ONE
    table whose column has several chunks (the reference chunks too,
    _internal/arrow_ops/transform_pyarrow.py:654-737); the limit is lowered here so that a
    small block crosses it."""
    from cluster_anywhere_b200 import device as dev

    monkeypatch.setattr(dev, "_OFFSET_LIMIT", 90 * 100 + 50)
    rng = np.random.default_rng(12)
    n = 1000
    pay = rng.integers(0, 256, n * 90, dtype=np.uint8)
    offs = np.arange(n + 1, dtype=np.int32) * 90
    t = pa.table({"k": pa.array(rng.integers(0, 1000, n), pa.int64()),
                  "p": pa.Array.from_buffers(pa.binary(), n, [None, pa.py_buffer(offs), pa.py_buffer(pay)])})
    rt = dev.default_runtime()
    dt = dev.stage_table(rt, t)
    out = dev.export_blocks(rt, dt, [(0, 350), (350, 350), (350, 1000)])
    assert [o.num_rows for o in out] == [350, 0, 650]
    assert out[0]["p"].num_chunks == 4 and out[2]["p"].num_chunks == 7  # 100 rows per chunk
    assert pa.concat_tables(out).equals(t) or pa.concat_tables(out).combine_chunks().equals(t.combine_chunks())

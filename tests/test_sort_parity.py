"""Sort path parity: B200 engine vs the oracle / the reference's golden vectors.

Every test body runs twice: against the host-memory double of the C-ABI (checks
the Python host layer here, without a GPU) and -- marked ``gpu`` -- against the
real sm_100a kernels through the C-ABI.  Bar: bit-exact rows (key order, tie
order, payload bytes, validity) and identical partition cut points.
"""
import random

import numpy as np
import pyarrow as pa
import pytest

import _golden as G
from cluster_anywhere_b200 import B200ArrowBlockAccessor, SortKey
from cluster_anywhere_b200.exchange import PullBasedShuffleTaskScheduler, SortTaskSpec
from oracle import exchange_oracle as O


def _check_partitions(got, exp):
    assert len(got) == len(exp)
    for g, e in zip(got, exp):
        assert G.tables_equal(g, e), (g.num_rows, e.num_rows)


# ---- the reference's KATs (python/ray/data/tests/test_util.py:305-377) --------

def test_kat_find_partitions(engine):
    def parts(table, boundaries, cols, desc):
        acc = B200ArrowBlockAccessor.for_block(table)
        return acc.sort_and_partition(boundaries, SortKey(cols, desc))

    p = parts(pa.table({"value": [1, 2, 2, 3, 5]}), [(0,), (2,), (4,), (6,)], ["value"], [False])
    assert [x.to_pydict().get("value", []) for x in p] == [[], [1], [2, 2, 3], [5], []]
    p = parts(pa.table({"value": [5, 3, 2, 2, 1]}), [(6,), (3,), (2,), (0,)], ["value"], [True])
    assert [x.to_pydict().get("value", []) for x in p] == [[], [5, 3], [2, 2], [1], []]
    t = pa.table({"col1": [1, 1, 1, 1, 1, 2, 2], "col2": [4, 3, 2.5, 2, 1, 2, 1]})
    p = parts(t, [(1, 3), (1, 2), (2, 2), (2, 0)], ["col1", "col2"], [False, True])
    assert [x.to_pydict().get("col2", []) for x in p] == [[4], [3, 2.5], [2, 1], [2, 1], []]
    t = pa.table({"col1": [2, 2, 1, 1, 1, 1, 1], "col2": [1, 2, 1, 2, 3, 4, 5]})
    p = parts(t, [(2, 0), (2, 2), (1, 2), (1, 6)], ["col1", "col2"], [True, False])
    assert [x.to_pydict().get("col2", []) for x in p] == [[], [1, 2], [1, 2], [3, 4, 5], []]
    p = parts(pa.table({"value": [1, 2, 3, None, None]}), [(2,), (4,)], ["value"], [False])
    assert [x.to_pydict().get("value", []) for x in p] == [[1], [2, 3], [None, None]]
    p = parts(pa.table({"value": [2, 2, 2, 2, 2]}), [(1,), (2,), (3,)], ["value"], [False])
    assert [x.to_pydict().get("value", []) for x in p] == [[], [], [2, 2, 2, 2, 2], []]
    # test_util.py:235-302 as cut positions (None boundary sorts after everything)
    p = parts(pa.table({"value": [1, 2, 3, None, None]}), [(None,)], ["value"], [False])
    assert [x.num_rows for x in p] == [5, 0]
    p = parts(pa.table({"value": [2, 2, 2, 2, 2]}), [(2,), (1,)], ["value"], [True])
    assert [x.num_rows for x in p] == [5, 0, 0]
    # None boundary components (NULL_SENTINEL, util.py:67-89): after every row on an
    # ascending column, BEFORE every row on a descending one (reversed search,
    # util.py:775-799).  Row counts taken from the reference itself (slices may overlap
    # when the cuts are not monotone -- util.py:823-827 slices regardless).
    for vals in ([5, 3, 2, 2, 1], [5, 3, None, 2, 1, None]):
        n = len(vals)
        t = pa.table({"value": pa.array(vals, pa.int64())})
        assert [x.num_rows for x in parts(t, [(None,)], ["value"], [True])] == [0, n]
        assert [x.num_rows for x in parts(t, [(3,), (None,)], ["value"], [True])] == [2, 0, n]
        assert [x.num_rows for x in parts(t, [(None,), (3,)], ["value"], [True])] == [0, 2, n - 2]
        assert [x.num_rows for x in parts(t, [(None,), (3,)], ["value"], [False])] == [n, 0, 2 * n - 8]
    t = pa.table({"k": [0, 0, 0, 0, 1, 1], "g": [2.0, 1.0, 0.0, -1.0, 5.0, 4.0]})
    assert [x.num_rows for x in parts(t, [(0, None)], ["k", "g"], [False, True])] == [0, 6]
    assert [x.num_rows for x in parts(t, [(0, None)], ["k", "g"], [False, False])] == [4, 2]
    assert [x.num_rows for x in parts(t, [(0, None)], ["k", "g"], [True, True])] == [2, 4]


def test_kat_seeded_two_key_sort(engine):
    # python/ray/data/tests/test_sort.py:165-183
    rnd = random.Random(0xDEED)
    rows = [{"a": rnd.randint(0, 9), "b": rnd.randint(0, 9)} for _ in range(1000)]
    t = pa.Table.from_pylist(rows)
    for desc in (False, True):
        s = B200ArrowBlockAccessor.for_block(t).sort(SortKey(["a", "b"], [desc, desc]))
        exp = sorted(((r["a"], r["b"]) for r in rows), reverse=desc)
        assert list(zip(s["a"].to_pylist(), s["b"].to_pylist())) == exp


def test_empty_block_contract(engine):
    # python/ray/data/tests/test_sort.py:186-234
    e = pa.table({"a": pa.array([], pa.int64())})
    sk = SortKey("a")
    acc = B200ArrowBlockAccessor.for_block(e)
    assert acc.sort(sk).num_rows == 0
    parts = acc.sort_and_partition([(1,), (2,)], sk)
    assert len(parts) == 3 and all(p.num_rows == 0 and p.num_columns == 0 for p in parts)
    merged, meta = B200ArrowBlockAccessor.merge_sorted_blocks([e, e], sk)
    assert merged.num_rows == 0 and meta.num_rows == 0
    assert acc.sample(10, sk).num_rows == 0
    assert SortTaskSpec.sample_boundaries([e, e], sk, 4) == [None, None, None]
    t = pa.table({"a": [3, 1, 2]})
    merged, _ = B200ArrowBlockAccessor.merge_sorted_blocks([e, t, e], sk)
    assert merged["a"].to_pylist() == [1, 2, 3]


def test_sort_key_errors(engine):
    with pytest.raises(ValueError):
        SortKey(key=5)
    with pytest.raises(ValueError):
        SortKey(key=["a", "b"], descending=[True])
    with pytest.raises(ValueError):
        SortKey(key="a", boundaries=["x"])
    t = pa.table({"a": [1]})
    with pytest.raises(ValueError):
        B200ArrowBlockAccessor.for_block(t).sort(SortKey("nope"))
    with pytest.raises(NotImplementedError):  # nested key types are not offloaded
        B200ArrowBlockAccessor.for_block(pa.table({"s": [[2], [1]]})).sort(SortKey("s"))


# ---- reference-generated golden fixtures ---------------------------------------

@pytest.mark.parametrize("case", G.cases("sort"), ids=lambda c: c["name"])
def test_golden_sort_map(engine, case):
    b = G.boundaries(case)
    sk = SortKey(case["columns"], case["descending"])
    for i, name in enumerate(case["inputs"]):
        out = SortTaskSpec.map(i, G.load(name), len(b) + 1, b, sk)
        parts, meta = out[:-1], out[-1]
        assert len(parts) == len(case["map_outputs"][i])
        for p, gold in zip(parts, case["map_outputs"][i]):
            assert G.tables_equal(p, G.load(gold)), gold
        assert meta.num_rows == G.load(name).num_rows


@pytest.mark.parametrize("case", G.cases("sort"), ids=lambda c: c["name"])
def test_golden_sort_exchange(engine, case):
    b = G.boundaries(case)
    sk = SortKey(case["columns"], case["descending"])
    blocks = [G.load(n) for n in case["inputs"]]
    spec = SortTaskSpec(b, sk)
    outs, stats = PullBasedShuffleTaskScheduler(spec).execute(blocks, len(b) + 1)
    assert set(stats) == {"map", "reduce"}
    for o, gold in zip(outs, case["outputs"]):
        assert G.tables_equal(o, G.load(gold)), gold


# ---- seeded random inputs vs the oracle -----------------------------------------

_DTYPES = ["int8", "int16", "int32", "int64", "uint8", "uint16", "uint32", "uint64",
           "float32", "float64"]


def _rand_col(rng, n, dtype, nulls, lo_card):
    if dtype.startswith("float"):
        v = rng.standard_normal(n).astype(dtype)
        if lo_card:
            v = np.round(v * 2).astype(dtype)
        v[rng.random(n) < 0.05] = np.nan
        v[rng.random(n) < 0.02] = np.inf
        v[rng.random(n) < 0.02] = -np.inf
        v[rng.random(n) < 0.03] = -0.0
        v[rng.random(n) < 0.03] = 0.0
    else:
        info = np.iinfo(dtype)
        if lo_card:
            v = rng.integers(max(info.min, -6), min(info.max, 6), n, endpoint=True).astype(dtype)
        else:
            v = rng.integers(info.min, info.max, n, dtype=dtype, endpoint=True)
            v[rng.random(n) < 0.01] = info.max
            v[rng.random(n) < 0.01] = info.min
    mask = (rng.random(n) < 0.15) if nulls else None
    return pa.array(v, mask=mask)


@pytest.mark.parametrize("dtype", _DTYPES)
@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("nulls", [False, True])
def test_single_key_all_types(engine, dtype, desc, nulls):
    rng = np.random.default_rng(hash((dtype, desc, nulls)) % (2**32))
    n = 20011 if engine == "b200" else 1500
    for lo_card in (False, True):
        t = pa.table({
            "k": _rand_col(rng, n, dtype, nulls, lo_card),
            "row": pa.array(np.arange(n, dtype=np.int32)),
            "s": pa.array([("x" * int(l)) for l in rng.integers(0, 9, n)]),
        })
        got = B200ArrowBlockAccessor.for_block(t).sort(SortKey("k", desc))
        exp = O.sort_table(t, ["k"], [desc])
        assert G.tables_equal(got, exp)


@pytest.mark.parametrize("trial", range(6))
def test_multi_key_with_boundaries(engine, trial):
    rng = np.random.default_rng(100 + trial)
    n = 30000 if engine == "b200" else 2000
    dts = [_DTYPES[(trial * 3 + j) % len(_DTYPES)] for j in range(3)]
    desc = [bool((trial >> j) & 1) for j in range(3)]
    t = pa.table({
        "a": _rand_col(rng, n, dts[0], trial % 2 == 1, True),
        "b": _rand_col(rng, n, dts[1], False, True),
        "c": _rand_col(rng, n, dts[2], trial % 3 == 2, False),
        "payload": pa.array([rng.bytes(int(l)) for l in rng.integers(0, 50, n)], type=pa.binary()),
    })
    cols = ["a", "b", "c"]
    # boundaries drawn from the data (as sampled boundaries are), ordered by the key
    srt = O.sort_table(t, cols, desc)
    pick = sorted(rng.integers(0, n, 5).tolist())
    bounds = []
    for p in pick:
        tup = tuple(srt[c][p].as_py() for c in cols[: 1 + trial % 3])
        if any(x is None or (isinstance(x, float) and x != x) for x in tup):
            continue
        bounds.append(tup)
    got = B200ArrowBlockAccessor.for_block(t).sort_and_partition(bounds, SortKey(cols, desc))
    exp = O.sort_and_partition(t, bounds, cols, desc)
    _check_partitions(got, exp)


def test_boundary_edge_values(engine):
    # non-representable / out-of-range boundaries (host normalisation in keys.py)
    t = pa.table({"i": pa.array([-128, -3, 0, 0, 5, 127], pa.int8()),
                  "f": pa.array([-1.5, 0.1, 0.1, 0.5, 3.0, np.inf], pa.float32())})
    for desc in (False, True):
        for col, bs in (("i", [-1000, -128, -2.5, 0, 0.5, 127, 128, 1e9, float("inf"), float("-inf")]),
                        ("f", [-1e39, -1.5, 0.1, 0.25, 0.5, 1e39, float("inf"), 3])):
            for b in bs:
                got = B200ArrowBlockAccessor.for_block(t).sort_and_partition([(b,)], SortKey(col, desc))
                exp = O.sort_and_partition(t, [(b,)], [col], [desc])
                assert [g.num_rows for g in got] == [e.num_rows for e in exp], (col, b, desc)


def test_graysort_rows(engine):
    # BASELINE.json configs[1] shape at parity size: int64 key + 90-byte binary payload
    rng = np.random.default_rng(2)
    n = 100_000 if engine == "b200" else 3000
    keys = rng.integers(0, 2**63 - 1, n, dtype=np.int64)
    payload = rng.integers(0, 256, n * 90, dtype=np.uint8)
    offs = np.arange(n + 1, dtype=np.int32) * 90
    c1 = pa.Array.from_buffers(pa.binary(), n, [None, pa.py_buffer(offs), pa.py_buffer(payload)])
    t = pa.table({"c_0": keys, "c_1": c1})
    q = np.quantile(keys.astype(np.float64), [0.25, 0.5, 0.75])
    bounds = [(int(x),) for x in q]
    for desc in (False, True):
        b = bounds[::-1] if desc else bounds
        got = B200ArrowBlockAccessor.for_block(t).sort_and_partition(b, SortKey("c_0", desc))
        exp = O.sort_and_partition(t, b, ["c_0"], [desc])
        _check_partitions(got, exp)


def test_config1_range_sort_descending(engine):
    # BASELINE.json configs[0]: range(N).sort('id', descending=True) over several blocks
    n, m = (400_000, 8) if engine == "b200" else (4000, 4)
    ids = np.arange(n, dtype=np.int64)
    blocks = [pa.table({"id": ids[i * n // m:(i + 1) * n // m]}) for i in range(m)]
    sk = SortKey("id", True)
    bounds = [((m - 1 - i) * n // m,) for i in range(m - 1)]
    outs, _ = PullBasedShuffleTaskScheduler(SortTaskSpec(bounds, sk)).execute(blocks, m)
    exp = O.sort_exchange(blocks, bounds, ["id"], [True])
    _check_partitions(outs, exp)
    allrows = pa.concat_tables(outs)["id"].to_numpy()
    assert (allrows == ids[::-1]).all()


def test_wide_and_chunked_payload_columns(engine):
    rng = np.random.default_rng(9)
    n = 5000 if engine == "b200" else 600
    k = rng.integers(0, 50, n)
    tens = pa.FixedSizeListArray.from_arrays(pa.array(rng.integers(0, 255, n * 7, dtype=np.uint8)), 7)
    ll = pa.array([rng.integers(0, 100, int(l)).astype(np.int16) for l in rng.integers(0, 6, n)],
                  type=pa.large_list(pa.int16()))
    t = pa.table({
        "k": k,
        "b": pa.array(rng.random(n) < 0.5, mask=rng.random(n) < 0.1),
        "fsb": pa.array([rng.bytes(5) for _ in range(n)], type=pa.binary(5)),
        "t7": tens,
        "ll": ll,
        "ls": pa.array([str(x) * 3 for x in k], type=pa.large_string()),
        "ts": pa.array(rng.integers(0, 10**12, n), type=pa.timestamp("us")),
    })
    # chunked input with a sliced (offset != 0) chunk
    t2 = pa.concat_tables([t.slice(0, n // 3), t.slice(n // 3, n // 2), t.slice(n // 3 + n // 2)])
    for table in (t, t2.slice(7)):
        got = B200ArrowBlockAccessor.for_block(table).sort(SortKey(["k", "ts"], [True, False]))
        exp = O.sort_table(table, ["k", "ts"], [True, False])
        assert G.tables_equal(got, exp.combine_chunks())


def test_take_and_sample(engine):
    rng = np.random.default_rng(3)
    n = 1000
    t = pa.table({"k": rng.integers(0, 9, n), "s": pa.array([str(i) for i in range(n)])})
    idx = rng.integers(0, n, 300)
    got = B200ArrowBlockAccessor.for_block(t).take(idx)
    assert G.tables_equal(got, t.take(idx))
    s = B200ArrowBlockAccessor.for_block(t).sample(50, SortKey("k"))
    assert s.num_rows == 50 and s.schema.names == ["k"]
    with pytest.raises(IndexError):
        B200ArrowBlockAccessor.for_block(t).take([n])


# ---- string / binary sort keys (BASELINE configs[4]) ----------------------------------

def _string_table(rng, n, kind="string", nulls=True, zipf=False):
    if zipf:
        vocab = ["w%d" % i + "x" * int(rng.integers(0, 24)) for i in range(200)]
        vals = [vocab[min(int(z), len(vocab)) - 1] for z in rng.zipf(1.2, n)]
    else:
        alphabet = ["", "a", "b", "ab", "aa", "B", "\u00e9", "abc", "ab\x00", "ab\x00c", "zzzzzzzz", "zzzzzzzzz",
                    "zzzzzzzz\x00", "0123456789abcdef", "0123456789abcdefg", "0123456789abcdeF"]
        vals = [alphabet[int(i)] + ("" if rng.random() < 0.7 else str(rng.integers(0, 50)))
                for i in rng.integers(0, len(alphabet), n)]
    mask = (rng.random(n) < 0.1) if nulls else None
    if kind == "string":
        key = pa.array(vals, pa.string(), mask=mask)
    elif kind == "large_string":
        key = pa.array(vals, pa.large_string(), mask=mask)
    else:
        key = pa.array([v.encode() + bytes([int(rng.integers(0, 256))]) for v in vals], pa.binary(), mask=mask)
    return pa.table({"s": key, "row": np.arange(n), "p": pa.array(["p%d" % i for i in range(n)])})


@pytest.mark.parametrize("kind", ["string", "large_string", "binary"])
@pytest.mark.parametrize("desc", [False, True])
def test_string_key_sort(engine, kind, desc):
    rng = np.random.default_rng(abs(hash((kind, desc))) % (2**32))
    n = 20_011 if engine == "b200" else 1200
    for zipf in (False, True):
        t = _string_table(rng, n, kind, nulls=True, zipf=zipf)
        got = B200ArrowBlockAccessor.for_block(t).sort(SortKey("s", desc))
        assert G.tables_equal(got, O.sort_table(t, ["s"], [desc]))
    # string as the SECOND key, numeric first
    t = _string_table(rng, n, kind).append_column("k", pa.array(rng.integers(0, 5, n)))
    got = B200ArrowBlockAccessor.for_block(t).sort(SortKey(["k", "s"], [desc, not desc]))
    assert G.tables_equal(got, O.sort_table(t, ["k", "s"], [desc, not desc]))
    # all-null / all-empty / empty-string-only columns
    t = pa.table({"s": pa.array([None, None, None], pa.string()), "row": [0, 1, 2]})
    assert G.tables_equal(B200ArrowBlockAccessor.for_block(t).sort(SortKey("s", desc)), O.sort_table(t, ["s"], [desc]))
    t = pa.table({"s": pa.array(["", "", None, ""], pa.string()), "row": [0, 1, 2, 3]})
    assert G.tables_equal(B200ArrowBlockAccessor.for_block(t).sort(SortKey("s", desc)), O.sort_table(t, ["s"], [desc]))


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("nulls", [False, True])
def test_string_key_partitions(engine, desc, nulls):
    """sort_and_partition / the pull-based exchange with a string key vs the oracle
    (find_partition_index on an object column: nulls are NULL_SENTINEL, util.py:765-772)."""
    rng = np.random.default_rng(77 + int(desc) * 2 + int(nulls))
    n = 9_001 if engine == "b200" else 700
    t = _string_table(rng, n, "string", nulls=nulls)
    bounds = [("B",), ("aa",), ("ab",), ("ab\x00c",), ("zzzzzzzz",)]
    if desc:
        bounds = bounds[::-1]
    if not (desc and nulls):
        bounds = bounds + [(None,)] if not desc else [(None,)] + bounds
    got = B200ArrowBlockAccessor.for_block(t).sort_and_partition(bounds, SortKey("s", desc))
    exp = O.sort_and_partition(t, bounds, ["s"], [desc])
    _check_partitions(got, exp)
    blocks = [t.slice(0, n // 3), t.slice(n // 3, 0), t.slice(n // 3)]
    from cluster_anywhere_b200.exchange import DeviceExchange

    out = DeviceExchange().sort_blocks(blocks, SortKey("s", desc), boundaries=bounds)
    exp_x = O.sort_exchange(blocks, bounds, ["s"], [desc])
    if not (desc and nulls):  # per-block vs global cuts agree wherever the rule is per-row consistent
        _check_partitions(out, exp_x)
    assert G.tables_equal(pa.concat_tables(out), O.sort_table(t, ["s"], [desc]))
    # sampled boundaries: whole exchange
    out = DeviceExchange().sort_blocks(blocks, SortKey("s", desc), seed=4)
    assert len(out) == 3
    assert G.tables_equal(pa.concat_tables([o for o in out if o.num_rows]), O.sort_table(t, ["s"], [desc]))

"""Pins oracle/exchange_oracle.py (CPU restatement) before anything trusts it.

1. The reference's own known-answer vectors for this path, restated from
   python/ray/data/tests/test_util.py:235-377 and test_sort.py:165-183.
2. Fixtures under tests/golden/ produced by running the reference itself
   (tests/golden/make_golden.py).
3. When /root/reference is present (build container only): live side-by-side.
"""
import random

import numpy as np
import pyarrow as pa
import pytest

import _golden as G
from oracle import exchange_oracle as O
from oracle import ref_shim


# ---- 1. reference KATs (test_util.py:235-377) --------------------------------

def test_kat_single_column_ascending():
    t = pa.table({"value": [1, 2, 2, 3, 5]})
    got = [O.find_partition_index(t, (v,), ["value"], [False]) for v in (0, 2, 4, 6)]
    assert got == [0, 1, 4, 5]


def test_kat_single_column_descending():
    t = pa.table({"value": [5, 3, 2, 2, 1]})
    got = [O.find_partition_index(t, (v,), ["value"], [True]) for v in (6, 3, 2, 0)]
    assert got == [0, 2, 4, 5]


def test_kat_multi_column():
    t = pa.table({"col1": [1, 1, 1, 2, 2], "col2": [3, 2, 1, 2, 1]})
    f = lambda d: O.find_partition_index(t, d, ["col1", "col2"], [False, True])
    assert [f((1, 3)), f((1, 2)), f((2, 2)), f((0, 4)), f((2, 0))] == [0, 1, 3, 0, 5]


def test_kat_nulls():
    t = pa.table({"value": [1, 2, 3, None, None]})
    f = lambda d: O.find_partition_index(t, d, ["value"], [False])
    assert [f((2,)), f((4,)), f((None,))] == [1, 3, 5]


def test_kat_duplicates():
    t = pa.table({"value": [2, 2, 2, 2, 2]})
    assert [O.find_partition_index(t, (v,), ["value"], [False]) for v in (2, 1, 3)] == [0, 0, 5]
    assert [O.find_partition_index(t, (v,), ["value"], [True]) for v in (2, 1, 3)] == [5, 5, 0]


def test_kat_find_partitions():
    t = pa.table({"value": [1, 2, 2, 3, 5]})
    p = O.find_partitions(t, [(0,), (2,), (4,), (6,)], ["value"], [False])
    assert [x.to_pydict()["value"] for x in p] == [[], [1], [2, 2, 3], [5], []]
    t = pa.table({"value": [5, 3, 2, 2, 1]})
    p = O.find_partitions(t, [(6,), (3,), (2,), (0,)], ["value"], [True])
    assert [x.to_pydict()["value"] for x in p] == [[], [5, 3], [2, 2], [1], []]
    t = pa.table({"col1": [1, 1, 1, 1, 1, 2, 2], "col2": [4, 3, 2.5, 2, 1, 2, 1]})
    p = O.find_partitions(t, [(1, 3), (1, 2), (2, 2), (2, 0)], ["col1", "col2"], [False, True])
    assert [x.to_pydict()["col2"] for x in p] == [[4], [3, 2.5], [2, 1], [2, 1], []]
    t = pa.table({"col1": [2, 2, 1, 1, 1, 1, 1], "col2": [1, 2, 1, 2, 3, 4, 5]})
    p = O.find_partitions(t, [(2, 0), (2, 2), (1, 2), (1, 6)], ["col1", "col2"], [True, False])
    assert [x.to_pydict()["col2"] for x in p] == [[], [1, 2], [1, 2], [3, 4, 5], []]
    t = pa.table({"value": [1, 2, 3, None, None]})
    p = O.find_partitions(t, [(2,), (4,)], ["value"], [False])
    assert [x.to_pydict()["value"] for x in p] == [[1], [2, 3], [None, None]]
    t = pa.table({"value": [2, 2, 2, 2, 2]})
    p = O.find_partitions(t, [(1,), (2,), (3,)], ["value"], [False])
    assert [x.to_pydict()["value"] for x in p] == [[], [], [2, 2, 2, 2, 2], []]


def test_kat_seeded_two_key_sort():
    # test_sort.py:165-183: Random(0xDEED), a in [0,10), b in [0,10); sort == sorted()
    rnd = random.Random(0xDEED)
    rows = [{"a": rnd.randint(0, 9), "b": rnd.randint(0, 9)} for _ in range(1000)]
    t = pa.Table.from_pylist(rows)
    for desc in (False, True):
        s = O.sort_table(t, ["a", "b"], [desc, desc])
        exp = sorted(((r["a"], r["b"]) for r in rows), reverse=desc)
        assert list(zip(s["a"].to_pylist(), s["b"].to_pylist())) == exp


def test_empty_block_contract():
    # test_sort.py:186-234
    e = pa.table({"a": pa.array([], pa.int64())})
    parts = O.sort_and_partition(e, [(1,), (2,)], ["a"], [False])
    assert len(parts) == 3 and all(p.num_rows == 0 and p.num_columns == 0 for p in parts)
    assert O.merge_sorted_blocks([e, e], ["a"], [False]).num_rows == 0
    assert O.boundaries_from_samples([], [False], 4) == [None, None, None]


def test_sort_is_stable_and_nulls_last_both_directions():
    t = pa.table({"k": pa.array([3, None, 1, float("nan"), 2, None, float("nan")]),
                  "i": list(range(7))})
    assert O.sort_table(t, ["k"], [False])["i"].to_pylist() == [2, 4, 0, 3, 6, 1, 5]
    assert O.sort_table(t, ["k"], [True])["i"].to_pylist() == [0, 4, 2, 3, 6, 1, 5]


# ---- 2. reference-generated fixtures ----------------------------------------

@pytest.mark.parametrize("case", G.cases("sort"), ids=lambda c: c["name"])
def test_golden_sort(case):
    blocks = [G.load(n) for n in case["inputs"]]
    b = G.boundaries(case)
    for i, blk in enumerate(blocks):
        parts = O.sort_and_partition(blk, b, case["columns"], case["descending"])
        assert len(parts) == len(case["map_outputs"][i])
        for p, name in zip(parts, case["map_outputs"][i]):
            assert G.tables_equal(p, G.load(name)), name
    outs = O.sort_exchange(blocks, b, case["columns"], case["descending"])
    for o, name in zip(outs, case["outputs"]):
        assert G.tables_equal(o, G.load(name)), name


@pytest.mark.parametrize("case", G.cases("shuffle"), ids=lambda c: c["name"])
def test_golden_shuffle(case):
    blocks = [G.load(n) for n in case["inputs"]]
    outs = O.shuffle_exchange(blocks, case["num_outputs"], case["seed"])
    for o, name in zip(outs, case["outputs"]):
        assert G.tables_equal(o, G.load(name)), name


@pytest.mark.parametrize("case", G.cases("aggregate"), ids=lambda c: c["name"])
def test_golden_aggregate(case):
    blocks = [G.load(n) for n in case["inputs"]]
    got = O.groupby_aggregate(blocks, case["key"], [tuple(a) for a in case["aggs"]])
    # (a block that holds only the null group has a null-typed key column in the reference:
    # the concatenation promotes it back)
    exp = pa.concat_tables([t for t in (G.load(n) for n in case["outputs"]) if t.num_rows],
                           promote_options="permissive")
    assert got.schema.names == exp.schema.names
    for n in exp.schema.names:
        assert got[n].type == exp[n].type, n
        if pa.types.is_string(exp[n].type) or pa.types.is_binary(exp[n].type):
            assert got[n].to_pylist() == exp[n].to_pylist(), n
            continue
        a, b = exp[n].to_numpy(), got[n].to_numpy()
        np.testing.assert_allclose(b, a, rtol=1e-12, atol=1e-12, equal_nan=True, err_msg=n)


# ---- 3. live reference (build container only) ---------------------------------

needs_ref = pytest.mark.skipif(not ref_shim.available(), reason="/root/reference absent")


@needs_ref
def test_live_reference_sort_random():
    R = ref_shim.load()
    rng = np.random.default_rng(5)
    for trial in range(4):
        blocks = [pa.table({"k": rng.integers(-50, 50, n), "f": rng.standard_normal(n),
                            "s": pa.array([str(x) for x in rng.integers(0, 30, n)])})
                  for n in (300, 0, 211)]
        blocks[1] = blocks[0].slice(0, 0)
        desc = [bool(trial & 1), bool(trial & 2)]
        cols = ["k", "s"]
        b = sorted([(int(rng.integers(-50, 50)), str(int(rng.integers(0, 30)))) for _ in range(3)],
                   reverse=desc[0])
        sk = R.SortKey(key=cols, descending=desc)
        maps = [R.SortTaskSpec.map(i, x, 4, b, sk) for i, x in enumerate(blocks)]
        ref = [R.SortTaskSpec.reduce(sk, "default", *[m[j] for m in maps])[0] for j in range(4)]
        mine = O.sort_exchange(blocks, b, cols, desc)
        for r, m in zip(ref, mine):
            assert G.tables_equal(r, m)


@needs_ref
def test_live_reference_boundaries_from_samples():
    R = ref_shim.load()
    rng = np.random.default_rng(6)
    samples = [(int(a), float(b)) for a, b in zip(rng.integers(0, 5, 40), rng.standard_normal(40))]
    samples[3] = (samples[3][0], float("nan"))
    # restate sort_task_spec.py:200-226 by running its own statements
    for desc in ([False, False], [True, False], [False, True]):
        exp = list(samples)
        for i, d in list(enumerate(desc))[::-1]:
            exp.sort(key=lambda s, i=i: R.NULL_SENTINEL if (s[i] is None or s[i] != s[i]) else s[i],
                     reverse=d)
        q = [int(x * (len(exp) - 1)) for x in np.linspace(0, 1, 6)]
        exp = [exp[i] for i in q[1:-1]]
        got = O.boundaries_from_samples(samples, desc, 5)
        assert [tuple(map(repr, g)) for g in got] == [tuple(map(repr, e)) for e in exp]


@needs_ref
@pytest.mark.parametrize("key", ["s", ["k", "s"], ["s", "k"], ["k", "f"]],
                         ids=lambda k: k if isinstance(k, str) else "+".join(k))
def test_live_reference_aggregate_general_keys(key):
    """String keys and several key columns: the oracle's groupby against the reference's own
    SortAggregateTaskSpec.map / reduce run side by side on random blocks (the float key has
    NaN but no nulls: see tests/golden/make_golden_groupby.py for the reference's behaviour
    when one key column holds both)."""
    R = ref_shim.load()
    A = R.aggregate
    rng = np.random.default_rng(abs(hash(str(key))) % (2**32))
    words = ["", "a", "ab", "b", "zz", "é", "B", "a" * 30]
    blocks = []
    for n in (250, 0, 180, 90):
        f = np.round(rng.standard_normal(n) * 2) / 2
        f[rng.random(n) < 0.1] = np.nan
        blocks.append(pa.table({
            "k": pa.array(rng.integers(-2, 3, n), pa.int32(), mask=rng.random(n) < 0.1),
            "s": pa.array([None if rng.random() < 0.1 else words[int(rng.integers(0, len(words)))]
                           for _ in range(n)], pa.string()),
            "f": pa.array(f, pa.float64()),
            "v": pa.array(rng.standard_normal(n), pa.float64(), mask=rng.random(n) < 0.1),
            "w": pa.array(rng.integers(-9, 9, n), pa.int64()),
        }))
    blocks[1] = blocks[0].slice(0, 0)
    aggs = [A.Count(), A.Sum("v"), A.Mean("v"), A.Min("w"), A.Max("v"), A.Std("v"), A.Sum("w")]
    spec = [("count",), ("sum", "v"), ("mean", "v"), ("min", "w"), ("max", "v"), ("std", "v", 1), ("sum", "w")]
    sk = R.SortKey(key=key)
    maps = [R.SortAggregateTaskSpec.map(i, b, 1, [], sk, aggs) for i, b in enumerate(blocks)]
    ref = R.SortAggregateTaskSpec.reduce(sk, aggs, "default", *[m[0] for m in maps])[0]
    got = O.groupby_aggregate(blocks, key, spec)
    assert got.schema.names == ref.schema.names
    for n in ref.schema.names:
        a, b = ref[n].to_pylist(), got[n].to_pylist()
        assert len(a) == len(b), n
        for x, y in zip(a, b):
            if isinstance(x, float) and isinstance(y, (float, int)):
                assert (x != x and y != y) or abs(x - y) <= 1e-9 * max(1.0, abs(x)), (n, x, y)
            else:
                assert x == y, (n, x, y)


@needs_ref
@pytest.mark.parametrize("ignore_nulls", [True, False])
def test_live_reference_abs_max(ignore_nulls):
    """AbsMax (python/ray/data/aggregate.py:330-354): the oracle against the reference's own
    map / reduce on int and float columns with nulls, NaN and groups that hold nothing else."""
    R = ref_shim.load()
    A = R.aggregate
    rng = np.random.default_rng(5 + int(ignore_nulls))
    blocks = []
    for n in (300, 0, 170):
        v = np.round(rng.standard_normal(n) * 4, 1)
        v[rng.random(n) < 0.1] = np.nan
        k = rng.integers(0, 12, n)
        vm = rng.random(n) < 0.15
        vm |= k == 3           # group 3: only nulls
        v[k == 4] = np.nan     # group 4: only NaN (and nulls)
        blocks.append(pa.table({
            "k": pa.array(k, pa.int32()),
            "v": pa.array(v, pa.float64(), mask=vm),
            "w": pa.array(rng.integers(-90, 60, n), pa.int64(), mask=rng.random(n) < 0.1),
            "u": pa.array(rng.integers(0, 200, n), pa.uint8()),
        }))
    aggs = [A.AbsMax("v", ignore_nulls=ignore_nulls), A.AbsMax("w", ignore_nulls=ignore_nulls),
            A.AbsMax("u", ignore_nulls=ignore_nulls)]
    spec = [("abs_max", "v"), ("abs_max", "w"), ("abs_max", "u")]
    sk = R.SortKey(key="k")
    maps = [R.SortAggregateTaskSpec.map(i, b, 1, [], sk, aggs) for i, b in enumerate(blocks)]
    ref = R.SortAggregateTaskSpec.reduce(sk, aggs, "default", *[m[0] for m in maps])[0]
    if not ignore_nulls:
        return  # the oracle restates the default policy only; the engine's test covers the other
    got = O.groupby_aggregate(blocks, "k", spec)
    assert got.schema.names == ref.schema.names
    for n in ref.schema.names:
        a, b = ref[n].to_pylist(), got[n].to_pylist()
        assert a == b or all((x is None and y is None) or (x is not None and y is not None and float(x) == float(y))
                             for x, y in zip(a, b)), (n, a, b)

"""User-level group-by surface: ``GroupedData.aggregate`` / ``_aggregate_on`` and the
shortcuts (python/ray/data/grouped_data.py:50-93, :270-560), ``_build_multicolumn_aggs``
(python/ray/data/dataset.py:5362-5388).  Expected values: oracle.groupby_aggregate (the
reference's map / reduce composition) and the reference's own known answers
(D/tests/test_all_to_all.py:391-449: 100 rows grouped by A % 3 count 34/33/33,
:452-524: sums 1683/1617/1650, :720-793: means 49.5/49.0/50.0)."""
import numpy as np
import pyarrow as pa
import pytest

from cluster_anywhere_b200 import Count, Max, Mean, Min, Std, Sum
from cluster_anywhere_b200 import planner as P
from cluster_anywhere_b200.grouped_data import GroupedData, build_multicolumn_aggs, groupby
from oracle import exchange_oracle as O
from test_aggregate import assert_tables_close


def _blocks(rng, sizes):
    out = []
    for n in sizes:
        out.append(pa.table({"k": pa.array(rng.integers(0, 7, n), pa.int32()),
                             "x": pa.array(rng.standard_normal(n), pa.float64()),
                             "y": pa.array(rng.integers(-100, 100, n), pa.int64())}))
    return out


def _cat(blocks):
    return pa.concat_tables([b for b in blocks if b.num_rows], promote_options="permissive")


def test_reference_known_answers_through_groupby(engine):
    xs = list(range(100))
    blocks = [pa.table({"A": [x % 3 for x in xs[i::4]], "B": xs[i::4]}) for i in range(4)]
    g = groupby(blocks, "A")
    assert _cat(g.count())["count()"].to_pylist() == [34, 33, 33]
    assert _cat(g.sum("B"))["sum(B)"].to_pylist() == [1683, 1617, 1650]
    assert _cat(g.mean("B"))["mean(B)"].to_pylist() == [49.5, 49.0, 50.0]
    assert _cat(g.min("B"))["min(B)"].to_pylist() == [0, 1, 2]
    assert _cat(g.max("B"))["max(B)"].to_pylist() == [99, 97, 98]


def test_aggregate_on_every_column_and_key_none(engine):
    rng = np.random.default_rng(8)
    blocks = _blocks(rng, [300, 0, 1200, 40])
    g = GroupedData(blocks, "k")
    got = _cat(g.sum())  # on=None: every non-key column (grouped_data.py:76-93)
    exp = O.groupby_aggregate(blocks, "k", [("sum", "x"), ("sum", "y")])
    assert got.schema.names == ["k", "sum(x)", "sum(y)"]
    assert_tables_close(got, exp, exact=("k",))
    got = _cat(g.aggregate(Count(), Mean("x"), Std("x", ddof=1), Min("y"), Max("y")))
    exp = O.groupby_aggregate(blocks, "k", [("count",), ("mean", "x"), ("std", "x", 1), ("min", "y"), ("max", "y")])
    assert_tables_close(got, exp, exact=("k",))
    # no key: one block, one row, no key column (planner/aggregate.py:61-63)
    out = GroupedData(blocks, None).aggregate(Sum("y"), Count())
    assert len(out) == 1 and out[0].num_rows == 1
    assert out[0]["sum(y)"].to_pylist() == [int(sum(b["y"].to_numpy().sum() for b in blocks))]
    assert out[0]["count()"].to_pylist() == [sum(b.num_rows for b in blocks)]
    # the push-based schedule gives the same blocks
    ctx = P.DataContext(shuffle_strategy=P.ShuffleStrategy.SORT_SHUFFLE_PUSH_BASED)
    got2 = _cat(GroupedData(blocks, "k", data_context=ctx).sum())
    assert_tables_close(got2, O.groupby_aggregate(blocks, "k", [("sum", "x"), ("sum", "y")]), exact=("k",))


def test_build_multicolumn_aggs():
    schema = pa.schema([("k", pa.int32()), ("a", pa.float64()), ("b", pa.float64())])
    aggs = build_multicolumn_aggs(schema, Sum, None, skip_cols="k", ignore_nulls=False)
    assert [a.name for a in aggs] == ["sum(a)", "sum(b)"]
    assert [a.name for a in build_multicolumn_aggs(schema, Mean, "a")] == ["mean(a)"]
    assert [a.name for a in build_multicolumn_aggs(schema, Max, ["b", "a"])] == ["max(b)", "max(a)"]
    with pytest.raises(ValueError):
        build_multicolumn_aggs(schema, Sum, [])
    with pytest.raises(ValueError):
        build_multicolumn_aggs(pa.schema([("k", pa.int32())]), Sum, None, skip_cols=["k"])
    with pytest.raises(ValueError):
        groupby([pa.table({"k": [1]})], "missing")

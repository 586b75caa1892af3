"""Pandas blocks: DataFrame in, DataFrame out (python/ray/data/_internal/pandas_block.py:
503-746).  The reference sorts a frame with ``sort_values`` (missing values last in both
directions) and merges with ``pd.concat`` + ``sort_values``; expectations here are pandas'
own stable ``sort_values`` of the same frames and the Arrow oracle on the converted rows."""
import numpy as np
import pandas as pd
import pyarrow as pa
import pytest

from cluster_anywhere_b200 import B200ArrowBlockAccessor, Count, Mean, SortKey, Sum
from cluster_anywhere_b200.block_accessor import B200PandasBlockAccessor
from cluster_anywhere_b200.exchange import PullBasedShuffleTaskScheduler
from cluster_anywhere_b200.exchange.aggregate_task_spec import SortAggregateTaskSpec
from cluster_anywhere_b200.exchange.shuffle_task_spec import ShuffleTaskSpec
from cluster_anywhere_b200.exchange.sort_task_spec import SortTaskSpec
from oracle import exchange_oracle as O


def _frame(rng, n):
    x = rng.standard_normal(n).round(1)
    x[rng.random(n) < 0.1] = np.nan
    return pd.DataFrame({"k": rng.integers(-5, 6, n), "x": x, "s": ["r%d" % i for i in range(n)]})


def _same(a: pd.DataFrame, b: pd.DataFrame):
    pd.testing.assert_frame_equal(a.reset_index(drop=True), b.reset_index(drop=True), check_dtype=False)


@pytest.mark.parametrize("desc", [False, True])
def test_sort_and_partition_frames(engine, desc):
    rng = np.random.default_rng(4)
    df = _frame(rng, 3000)
    acc = B200ArrowBlockAccessor.for_block(df)
    assert isinstance(acc, B200PandasBlockAccessor) and acc.to_block() is df
    out = acc.sort(SortKey("k", desc))
    assert isinstance(out, pd.DataFrame)
    _same(out, df.sort_values(by=["k"], ascending=[not desc], kind="stable"))
    # float key with missing values: last in both directions, like na_position="last"
    out = acc.sort(SortKey("x", desc))
    _same(out, df.sort_values(by=["x"], ascending=[not desc], kind="stable", na_position="last"))
    bounds = [(3,), (0,), (-2,)] if desc else [(-2,), (0,), (3,)]
    parts = acc.sort_and_partition(bounds, SortKey("k", desc))
    assert all(isinstance(p, pd.DataFrame) for p in parts) and len(parts) == 4
    exp = O.sort_and_partition(pa.Table.from_pandas(df, preserve_index=False), bounds, ["k"], [desc])
    for p, e in zip(parts, exp):
        assert len(p) == e.num_rows
        if e.num_rows:
            _same(p, e.to_pandas())
    empty = B200ArrowBlockAccessor.for_block(pd.DataFrame()).sort_and_partition(bounds, SortKey("k"))
    assert len(empty) == 4 and all(isinstance(p, pd.DataFrame) and len(p) == 0 for p in empty)


def test_sort_exchange_of_frames(engine):
    rng = np.random.default_rng(6)
    frames = [_frame(rng, n) for n in (700, 0, 1200, 90)]
    sk = SortKey("k")
    spec = SortTaskSpec([(-1,), (2,)], sk)
    out, stats = PullBasedShuffleTaskScheduler(spec).execute(frames, 3)
    assert len(out) == 3 and all(isinstance(b, pd.DataFrame) for b in out)
    whole = pd.concat([f for f in frames if len(f)], ignore_index=True)
    _same(pd.concat(out, ignore_index=True), whole.sort_values(by=["k"], kind="stable"))
    assert out[0]["k"].max() < -1 and out[1]["k"].between(-1, 1).all() and out[2]["k"].min() >= 2
    got, meta = B200PandasBlockAccessor.merge_sorted_blocks(
        [f.sort_values(by=["k"], kind="stable") for f in frames], sk)
    assert isinstance(got, pd.DataFrame) and meta.num_rows == len(whole)
    _same(got, whole.sort_values(by=["k"], kind="stable"))


def test_shuffle_and_aggregate_of_frames(engine):
    rng = np.random.default_rng(7)
    frames = [_frame(rng, n) for n in (500, 300, 900)]
    out, _ = PullBasedShuffleTaskScheduler(ShuffleTaskSpec(1 << 30, random_shuffle=True, random_seed=5)).execute(frames, 3)
    assert all(isinstance(b, pd.DataFrame) for b in out)
    whole = pd.concat(frames, ignore_index=True)
    assert sorted(pd.concat(out)["s"].tolist()) == sorted(whole["s"].tolist())
    spec = SortAggregateTaskSpec([(0,)], SortKey("k"), [Count(), Sum("x"), Mean("x")], batch_format="pandas")
    out, _ = PullBasedShuffleTaskScheduler(spec).execute(frames, 2)
    assert all(isinstance(b, pd.DataFrame) for b in out)
    got = pd.concat(out, ignore_index=True)
    exp = whole.groupby("k").agg(**{"count()": ("k", "size"), "sum(x)": ("x", "sum"), "mean(x)": ("x", "mean")}).reset_index()
    assert got["k"].tolist() == exp["k"].tolist() and got["count()"].tolist() == exp["count()"].tolist()
    np.testing.assert_allclose(got["sum(x)"].to_numpy(), exp["sum(x)"].to_numpy(), rtol=1e-6)   # fp64, 1e-6 relative
    np.testing.assert_allclose(got["mean(x)"].to_numpy(), exp["mean(x)"].to_numpy(), rtol=1e-6)

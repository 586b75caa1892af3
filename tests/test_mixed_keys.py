"""Range partitioning when the key TUPLE mixes numeric and string / binary columns
(xchg_partition_bounds_mixed on the sorted table, xchg_assign_partitions_mixed per row).

Reference: find_partition_index narrows [left, right) column by column whatever the column
types (python/ray/data/_internal/util.py:757-802): object columns get NULL_SENTINEL for
their nulls and for a None component, descending columns go through numpy.searchsorted with
a reversing sorter.  Oracle: oracle.sort_and_partition / find_partitions (that call
sequence); the reference itself made tests/golden/sortx_* (make_golden_mixed.py), which the
golden tests of test_sort_parity.py run.
"""
import itertools

import numpy as np
import pyarrow as pa
import pytest

from _golden import tables_equal
from cluster_anywhere_b200 import B200ArrowBlockAccessor, SortKey
from cluster_anywhere_b200 import device as dev
from cluster_anywhere_b200.exchange import DeviceExchange
from cluster_anywhere_b200.exchange.comm import run_virtual_world
from oracle import exchange_oracle as O

WORDS = [b"", b"\x00", b"a", b"a\x00", b"ab", b"abc", b"abcdefgh", b"abcdefghi", b"b", b"m", b"zz", b"\xff\xff"]


def _table(rng, n, s_nulls=True, k_nulls=False, typ=pa.binary()):
    s = [WORDS[int(i)] for i in rng.integers(0, len(WORDS), n)]
    if typ == pa.string():
        s = [w.decode("latin-1") for w in s]
    f = rng.integers(-1, 3, n).astype(np.float64)
    f[rng.random(n) < 0.05] = np.nan
    return pa.table({
        "k": pa.array(rng.integers(-2, 3, n), pa.int32(), mask=(rng.random(n) < 0.15) if k_nulls else None),
        "s": pa.array(s, typ, mask=(rng.random(n) < 0.2) if s_nulls else None),
        "f": pa.array(f),
        "row": pa.array(np.arange(n), pa.int64()),
    })


def _check(got, exp):
    assert len(got) == len(exp)
    for j, (g, e) in enumerate(zip(got, exp)):
        assert g.num_rows == e.num_rows, (j, [x.num_rows for x in got], [x.num_rows for x in exp])
        assert tables_equal(g, e), j


def _bounds(rng, cols, desc, nb):
    """Random boundary tuples (full, short, with None components), in sort order."""
    vals = {"k": [-3, -2, 0, 0.5, 1, 2, 7], "s": WORDS + [b"aa", b"zzz"], "f": [-1.0, 0.0, 0.5, 2.0]}
    out = []
    for _ in range(nb):
        ln = int(rng.integers(1, len(cols) + 1))
        b = []
        for c in cols[:ln]:
            v = vals[c][int(rng.integers(0, len(vals[c])))]
            if c == "s" and rng.random() < 0.15:
                v = None
            b.append(v)
        out.append(tuple(b))

    # the planner hands boundaries over in sort order (planner/sort.py:59-62): order the
    # tuples the way the rows are ordered
    class K:
        def __init__(self, t):
            self.t = t

        def __lt__(self, o):
            for a, b, d in zip(self.t, o.t, desc):
                if a == b or (a is None and b is None):
                    continue
                if a is None:
                    return False  # None: after every value, both directions (as the rows' nulls sort)
                if b is None:
                    return True
                return (a > b) if d else (a < b)
            return len(self.t) < len(o.t)

    return sorted(out, key=K)


@pytest.mark.parametrize("cols,desc", [
    (["k", "s"], [False, False]), (["s", "k"], [False, False]), (["k", "s"], [True, False]),
    (["f", "s", "k"], [False, False, True]), (["s", "f"], [False, True]), (["k", "s", "f"], [False, False, False]),
])
@pytest.mark.parametrize("k_nulls", [False, True])
def test_sort_and_partition_mixed_tuple_vs_oracle(engine, cols, desc, k_nulls):
    rng = np.random.default_rng(len(cols) * 7 + sum(desc) + int(k_nulls))
    n = 4000 if engine == "b200" else 500
    for trial in range(3):
        t = _table(rng, n, s_nulls=True, k_nulls=k_nulls)
        bounds = _bounds(rng, cols, desc, 6)
        got = B200ArrowBlockAccessor.for_block(t).sort_and_partition(bounds, SortKey(cols, desc))
        exp = O.sort_and_partition(t, bounds, cols, desc)
        _check(got, exp)


@pytest.mark.parametrize("cols,desc", [(["k", "s"], [False, True]), (["s", "k"], [True, False]),
                                       (["s", "k"], [True, True]), (["f", "s"], [True, True])])
def test_descending_string_column_replays_numpy_probes(engine, cols, desc):
    """Descending string columns: the reversed-view search of util.py:775-799, with and
    without nulls (nulls sit last in storage = first in the view, where numpy does not expect
    the greatest value; the outcome depends on the probe sequence and is replayed)."""
    rng = np.random.default_rng(91)
    for s_nulls in (False, True):
        t = _table(rng, 700, s_nulls=s_nulls, typ=pa.string())
        for bounds in ([(1, "m")], [("m", 1)], [("ab", 0)], [(0, "ab"), (0, "a")], [(None, 1)], [(0, None)],
                       [("zz",)], [(2.0, "b")], [(-1.0, "a\x00")]):
            b = [tuple(x) for x in bounds]
            # a component must be comparable with its column: skip the shapes that are not
            if any(isinstance(v, str) != (c == "s") for x in b for v, c in zip(x, cols) if v is not None):
                continue
            got = B200ArrowBlockAccessor.for_block(t).sort_and_partition(b, SortKey(cols, desc))
            exp = O.sort_and_partition(t, b, cols, desc)
            _check(got, exp)


def test_exchange_with_mixed_keys(engine):
    """The whole exchange (task specs, and the device-resident exchange) with explicit and
    with sampled boundaries."""
    from cluster_anywhere_b200.exchange import PullBasedShuffleTaskScheduler, SortTaskSpec

    rng = np.random.default_rng(12)
    t = _table(rng, 1500, s_nulls=True)
    blocks = [t.slice(0, 600), t.slice(600, 0), t.slice(600)]
    cols, desc = ["k", "s"], [False, False]
    bounds = [(-1, b"ab"), (0, b""), (0, b"m"), (1, None), (2, b"zz")]
    sk = SortKey(cols, desc)
    outs, _ = PullBasedShuffleTaskScheduler(SortTaskSpec(bounds, sk)).execute(blocks, len(bounds) + 1)
    _check(outs, O.sort_exchange(blocks, bounds, cols, desc))
    out = DeviceExchange().sort_blocks(blocks, sk, boundaries=bounds)
    _check(out, O.sort_exchange(blocks, bounds, cols, desc))
    out = DeviceExchange().sort_blocks(blocks, sk, seed=5)
    assert len(out) == 3
    assert tables_equal(pa.concat_tables([o for o in out if o.num_rows]), O.sort_table(t, cols, desc))


@pytest.mark.parametrize("cols,desc", [(["k", "s"], [False, False]), (["s", "k"], [False, True]),
                                       (["s", "f"], [True, False])])
def test_distributed_sort_with_mixed_keys(engine, cols, desc):
    """N ranks (virtual world): the map side assigns every row its destination with the
    per-row form of the rule; the result must be the global stable sort, cut at the
    boundaries like the reference cuts the merged rows."""
    rng = np.random.default_rng(3)
    t = _table(rng, 1800, s_nulls=not desc[cols.index("s")])
    per_rank = [[t.slice(0, 500), t.slice(500, 100)], [t.slice(600, 1200)], [t.slice(1800, 0)]]
    sk = SortKey(cols, desc)
    device = None
    if engine == "b200":
        import torch

        device = torch.device("cuda", torch.cuda.current_device())

    def run(boundaries):
        def fn(comm, r):
            ex = DeviceExchange(dev.Runtime())
            return ex.sort_blocks_distributed(per_rank[r], sk, boundaries, 9, group=comm)

        return [b for out in run_virtual_world(3, fn, device) for b in out]

    full_sorted = O.sort_table(t, cols, desc)
    got = run(None)
    assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), full_sorted)
    vals = {"k": 0, "s": b"ab", "f": 1.0}
    bounds = [tuple(vals[c] for c in cols)]
    got = run(bounds)
    _check(got, O.find_partitions(full_sorted, bounds, cols, desc))


def test_mixed_boundary_component_of_the_wrong_kind_raises(engine):
    t = _table(np.random.default_rng(1), 50)
    with pytest.raises(TypeError):
        B200ArrowBlockAccessor.for_block(t).sort_and_partition([(0, 5)], SortKey(["k", "s"]))

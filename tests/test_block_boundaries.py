"""Group boundaries of key-sorted blocks and distinct values on the device.

Known answers restated from the reference's python/ray/data/tests/test_block_boundaries.py
(:6-62: integer columns, NaN and None runs), plus random blocks against a numpy statement
of ``_get_block_boundaries`` (python/ray/data/block.py:515-592) and, in the build container,
the reference's own function run side by side."""
import numpy as np
import pyarrow as pa
import pytest

from cluster_anywhere_b200 import B200ArrowBlockAccessor
from cluster_anywhere_b200.aggregate import _get_block_boundaries
from cluster_anywhere_b200.exchange import DeviceExchange
from oracle import ref_shim


def test_groupby_map_groups_get_block_boundaries(engine):
    # tests/test_block_boundaries.py:6-27
    indices = _get_block_boundaries([np.array([1, 1, 2, 2, 3, 3]), np.array([1, 1, 2, 2, 3, 4]),
                                     np.array([1, 1, 2, 2, 3, 3])])
    assert list(indices) == [0, 2, 4, 5, 6]
    indices = _get_block_boundaries([np.array([1, 1, 2, 2, 3, 3]), np.array([1, 1, 2, 2, 3, 3]),
                                     np.array([1, 1, 2, 3, 3, 3])])
    assert list(indices) == [0, 2, 3, 4, 6]


def test_groupby_map_groups_get_block_boundaries_with_nan(engine):
    # tests/test_block_boundaries.py:30-62: NaN / None runs are one group each
    indices = _get_block_boundaries([np.array([1.0, 1.0, 2.0, 2.0, np.nan, np.nan]),
                                     np.array([1.0, 1.0, 2.0, 2.0, 3.0, 3.0])])
    assert list(indices) == [0, 2, 4, 6]
    indices = _get_block_boundaries([np.array([1.0, 1.0, 2.0, np.nan, np.nan, np.nan]),
                                     np.array(["a", "a", "b", "b", None, None], dtype=object)])
    assert list(indices) == [0, 2, 3, 4, 6]
    indices = _get_block_boundaries([np.array(["a", "a", None, None, None, None], dtype=object)])
    assert list(indices) == [0, 2, 6]


def _numpy_boundaries(table, keys):
    n = table.num_rows
    same = np.ones(n, bool)
    same[0] = False
    for k in keys:
        v = table[k].to_pylist()
        for i in range(1, n):
            a, b = v[i], v[i - 1]
            eq = (a is None and b is None) or (a is not None and b is not None and (
                a == b or (isinstance(a, float) and isinstance(b, float) and a != a and b != b)))
            same[i] &= eq
    return np.concatenate([np.flatnonzero(~same), [n]])


@pytest.mark.parametrize("keys", [["k"], ["s"], ["k", "s"], ["f", "k"], ["s", "f", "k"]], ids="+".join)
def test_block_boundaries_random(engine, keys):
    rng = np.random.default_rng(len("".join(keys)))
    n = 20_011 if engine == "b200" else 600
    f = np.round(rng.standard_normal(n))
    f[rng.random(n) < 0.1] = np.nan
    t = pa.table({
        "k": pa.array(rng.integers(0, 5, n), pa.int64(), mask=rng.random(n) < 0.1),
        "s": pa.array([None if rng.random() < 0.1 else "ab"[int(rng.integers(0, 2))] * int(rng.integers(0, 3))
                       for _ in range(n)]),
        "f": pa.array(f, pa.float64(), mask=rng.random(n) < 0.05),
    })
    t = t.sort_by([(k, "ascending") for k in keys])
    got = B200ArrowBlockAccessor.for_block(t).get_block_boundaries(keys)
    assert np.array_equal(got, _numpy_boundaries(t, keys))
    if ref_shim.available() and "f" not in keys:  # the reference function, live (build container)
        ref_shim.load()  # installs the import hooks for the reference's modules
        from ray.data.block import _get_block_boundaries as ref_fn

        cols = [np.array(t[k].to_pylist(), dtype=object) if pa.types.is_string(t[k].type)
                else t[k].to_numpy(zero_copy_only=False) for k in keys]
        assert np.array_equal(got, ref_fn(cols))
    assert list(B200ArrowBlockAccessor.for_block(t.slice(0, 0)).get_block_boundaries(keys)) == [0, 0]
    assert list(B200ArrowBlockAccessor.for_block(t).get_block_boundaries([])) == [0, n]


def test_unique(engine):
    rng = np.random.default_rng(3)
    n = 50_000 if engine == "b200" else 700
    blocks = []
    for m in (n, 0, n // 3):
        blocks.append(pa.table({
            "i": pa.array(rng.integers(-40, 40, m), pa.int32(), mask=rng.random(m) < 0.05),
            "s": pa.array([None if rng.random() < 0.05 else "w%d" % rng.integers(0, 60) for _ in range(m)], pa.string()),
            "f": pa.array(np.where(rng.random(m) < 0.1, np.nan, np.round(rng.standard_normal(m) * 4) / 4)),
        }))
    full = pa.concat_tables(blocks)
    ex = DeviceExchange()
    for col in ("i", "s"):
        got = ex.unique_blocks(blocks, col)
        vals = full[col].to_pylist()
        exp = sorted({v for v in vals if v is not None}) + ([None] if None in vals else [])
        assert got == exp
        assert B200ArrowBlockAccessor.for_block(full).unique(col) == exp
    got = ex.unique_blocks(blocks, "f")
    vals = full["f"].to_pylist()
    exp = sorted({v for v in vals if v == v})
    assert got[:len(exp)] == exp and len(got) == len(exp) + 1 and got[-1] != got[-1]  # one NaN, last
    assert ex.unique_blocks([], "i") == []

"""String / binary sort keys: the MSD walk (one sort by the first 8 bytes, then only the rows
whose order is still open, 8 bytes at a time, the length last) against pyarrow's order --
what ``pac.sort_indices`` gives the reference at
python/ray/data/_internal/arrow_ops/transform_pyarrow.py:41: bytewise (unsigned), a proper
prefix first, nulls last in both directions, ties in input order."""
import numpy as np
import pyarrow as pa
import pytest

from _golden import tables_equal
from cluster_anywhere_b200 import B200ArrowBlockAccessor, SortKey
from oracle import exchange_oracle as O


def _words(rng, n, kind):
    if kind == "zipf":          # heavy hitters: long runs of identical keys
        vocab = ["w%07d-%s" % (i, "x" * int(rng.integers(0, 20))) for i in range(200)]
        return [vocab[min(int(z), 200) - 1] for z in rng.zipf(1.3, n)]
    if kind == "prefix":        # long shared prefixes: the order is decided in chunk 2, 3, ...
        return ["https://example.org/a/very/long/common/prefix/" + "%d" % rng.integers(0, 5000) + "/" * int(rng.integers(0, 3))
                for _ in range(n)]
    if kind == "zeros":         # embedded and trailing zero bytes, proper prefixes, empty strings
        alphabet = [b"", b"\x00", b"a", b"a\x00", b"a\x00\x00", b"ab", b"abcdefgh", b"abcdefgh\x00", b"abcdefghi",
                    b"abcdefgh\x00\x00\x00\x00\x00\x00\x00\x00", b"\xff" * 8, b"\xff" * 9, b"\xff" * 17]
        return [alphabet[int(i)] for i in rng.integers(0, len(alphabet), n)]
    return ["".join(chr(int(c)) for c in rng.integers(97, 100, int(rng.integers(0, 30)))) for _ in range(n)]


@pytest.mark.parametrize("kind", ["zipf", "prefix", "zeros", "random"])
@pytest.mark.parametrize("desc", [False, True])
def test_string_key_sort(engine, kind, desc):
    rng = np.random.default_rng(31)
    n = 6000
    words = _words(rng, n, kind)
    mask = rng.random(n) < 0.08
    typ = pa.binary() if kind == "zeros" else pa.string()
    t = pa.table({"s": pa.array([None if m else w for w, m in zip(words, mask)], typ),
                  "i": pa.array(np.arange(n), pa.int64())})
    got = B200ArrowBlockAccessor.for_block(t).sort(SortKey("s", desc))
    assert tables_equal(got, O.sort_table(t, ["s"], [desc]))


def test_string_key_inside_a_multi_key_sort(engine):
    rng = np.random.default_rng(5)
    n = 5000
    t = pa.table({"a": pa.array(rng.integers(0, 4, n), pa.int32()),
                  "s": pa.array(_words(rng, n, "zipf")),
                  "b": pa.array(rng.integers(0, 3, n), pa.int16()),
                  "i": pa.array(np.arange(n), pa.int64())})
    for cols, desc in ((["a", "s"], [False, True]), (["s", "b"], [False, False]), (["a", "s", "b"], [True, False, True])):
        got = B200ArrowBlockAccessor.for_block(t).sort(SortKey(cols, desc))
        assert tables_equal(got, O.sort_table(t, cols, desc)), (cols, desc)


@pytest.mark.parametrize("desc", [False, True])
def test_primary_string_key_is_copied_from_one_row_per_run(engine, desc):
    """A sort whose primary key is a string column with repeated values hands out a second
    take index for that column -- one source row per run of identical keys -- and the column
    built from it is byte-identical to the plain take by the sort permutation."""
    import torch

    from cluster_anywhere_b200 import device as dev

    rng = np.random.default_rng(77)
    n = 5000
    words = _words(rng, n, "zipf")
    # two strings that share their first 8 bytes: their tie group stays open past level 0
    words[::50] = ["w0000001-tail-a" if i % 2 else "w0000001-tail-b" for i in range(len(words[::50]))]
    mask = rng.random(n) < 0.05
    t = pa.table({"s": pa.array([None if m else w for w, m in zip(words, mask)], pa.string()),
                  "i": pa.array(np.arange(n), pa.int64())})
    rt = dev.default_runtime()
    dt = dev.stage_table(rt, t)
    hints = {}
    perm, _ = dev.sort_permutation(rt, dt, ["s"], [desc], False, hints)
    assert "string_runs" in hints and hints["string_runs"][0] == "s"
    rows = hints["string_runs"][1]
    rt.synchronize()
    p, r = perm.cpu().numpy().view(np.uint32), rows.cpu().numpy().view(np.uint32)
    assert (p != r).any(), "repeated keys must share a source row"
    same = dev.export_table(rt, dev.take_table(rt, dt, perm, n, {"s": ("rows", rows)}))
    plain = dev.export_table(rt, dev.take_table(rt, dt, perm, n))
    assert tables_equal(same, plain) and tables_equal(same, O.sort_table(t, ["s"], [desc]))
    # the accessor path uses it by itself
    assert tables_equal(B200ArrowBlockAccessor.for_block(t).sort(SortKey("s", desc)), plain)
    # a secondary string key gets no hint (its runs are not runs of the final order)
    hints2 = {}
    dev.sort_permutation(rt, dt, ["i", "s"], [False, desc], False, hints2)
    assert "string_runs" not in hints2

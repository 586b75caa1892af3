"""Loader for the reference-generated fixtures under tests/golden/."""
import json
import os

import pyarrow as pa

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def manifest():
    with open(os.path.join(GOLDEN, "manifest.json")) as f:
        return json.load(f)


def cases(op):
    return [c for c in manifest()["cases"] if c["op"] == op]


def load(name) -> pa.Table:
    with pa.OSFile(os.path.join(GOLDEN, name), "rb") as f:
        return pa.ipc.open_file(f).read_all()


def boundaries(case):
    return [tuple(b) for b in case["boundaries"]]


def _column_bits_equal(x: pa.ChunkedArray, y: pa.ChunkedArray) -> bool:
    if x.type != y.type or len(x) != len(y):
        return False
    if pa.types.is_floating(x.type):
        # Arrow's equals() says NaN != NaN; compare validity and raw bit patterns.
        import numpy as np
        import pyarrow.compute as pac

        xv = pac.is_valid(x).to_numpy(zero_copy_only=False)
        yv = pac.is_valid(y).to_numpy(zero_copy_only=False)
        if not (xv == yv).all():
            return False
        xa = pac.fill_null(x, 0).to_numpy(zero_copy_only=False)
        ya = pac.fill_null(y, 0).to_numpy(zero_copy_only=False)
        u = {2: "u2", 4: "u4", 8: "u8"}[xa.dtype.itemsize]
        return bool((xa.view(u)[xv] == ya.view(u)[yv]).all())
    return x.equals(y)


def tables_equal(a: pa.Table, b: pa.Table) -> bool:
    """Bit-exact table equality; two empty tables are equal whatever their schema
    (the reference hands out schema-less empties, arrow_block.py:445-446)."""
    if a.num_rows == 0 and b.num_rows == 0:
        return True
    if a.schema.names != b.schema.names or a.num_rows != b.num_rows:
        return False
    return all(_column_bits_equal(a[n], b[n]) for n in a.schema.names)

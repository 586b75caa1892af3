"""Adds sort cases whose key tuple mixes numeric and string columns to
tests/golden/manifest.json by RUNNING THE REFERENCE's SortTaskSpec.map / reduce through
oracle/ref_shim.py (find_partition_index narrows column by column whatever the column
types, python/ray/data/_internal/util.py:757-802).  Run in the build container only
(needs /root/reference):

    python tests/golden/make_golden_mixed.py
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import pyarrow as pa

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_shim  # noqa: E402

WORDS = ["", "a", "ab", "abc", "b", "m", "mm", "zebra", "zz", "été"]


def _write(name: str, table: pa.Table) -> str:
    with pa.OSFile(os.path.join(HERE, name), "wb") as f:
        with pa.ipc.new_file(f, table.schema) as w:
            w.write_table(table)
    return name


def blocks_for(rng, sizes, s_nulls, k_nulls):
    out = []
    for n in sizes:
        s = [WORDS[int(i)] for i in rng.integers(0, len(WORDS), n)]
        out.append(pa.table({
            "k": pa.array(rng.integers(-2, 3, n, dtype=np.int64), mask=(rng.random(n) < 0.15) if k_nulls else None),
            "s": pa.array(s, pa.string(), mask=(rng.random(n) < 0.2) if s_nulls else None),
            "f": pa.array(rng.integers(0, 3, n).astype(np.float32)),
            "row": pa.array(np.arange(n, dtype=np.int32)),
        }))
    return out


def main():
    R = ref_shim.load()
    path = os.path.join(HERE, "manifest.json")
    with open(path) as fh:
        manifest = json.load(fh)
    manifest["cases"] = [c for c in manifest["cases"] if not c["name"].startswith("sortx_")]
    manifest["generator_mixed"] = "tests/golden/make_golden_mixed.py"

    def case(name, columns, descending, boundaries, seed, sizes=(260, 0, 190, 1), s_nulls=True, k_nulls=False):
        rng = np.random.default_rng(seed)
        blocks = blocks_for(rng, sizes, s_nulls, k_nulls)
        sk = R.SortKey(key=columns, descending=descending)
        nred = len(boundaries) + 1
        maps = [R.SortTaskSpec.map(i, b, nred, boundaries, sk) for i, b in enumerate(blocks)]
        outs = [R.SortTaskSpec.reduce(sk, "default", *[m[j] for m in maps])[0] for j in range(nred)]
        manifest["cases"].append({
            "name": name, "op": "sort", "columns": columns, "descending": descending,
            "boundaries": [list(b) for b in boundaries],
            "inputs": [_write(f"{name}.in{i}.arrow", b) for i, b in enumerate(blocks)],
            "map_outputs": [[_write(f"{name}.map{i}_{j}.arrow", m[j]) for j in range(nred)]
                            for i, m in enumerate(maps)],
            "outputs": [_write(f"{name}.out{j}.arrow", o) for j, o in enumerate(outs)],
        })

    case("sortx_int_string", ["k", "s"], [False, False], [(-1, "ab"), (0, ""), (0, "m"), (1, None), (2, "zz")], 41)
    case("sortx_string_int", ["s", "k"], [False, False], [("a", 1), ("m", -5), ("m", 2), (None, 0)], 42)
    case("sortx_int_string_desc", ["k", "s"], [False, True], [(-1, "m"), (0, "zz"), (0, "a"), (2, "")], 43,
         s_nulls=False)
    case("sortx_string_desc_first", ["s", "f"], [True, False], [("zebra", 1.0), ("m", 0.5), ("ab", 2.0)], 44,
         s_nulls=False)
    case("sortx_three_keys", ["f", "s", "k"], [False, False, True], [(0.0, "b", 1), (1.0, "ab"), (1.0, "ab", -1), (2.0,)],
         45)
    case("sortx_int_nulls_string", ["k", "s"], [False, False], [(-1, "b"), (1, "m"), (2, "zebra")], 46, k_nulls=True)
    with open(path, "w") as fh:
        json.dump(manifest, fh, indent=1)
    print("cases now:", len(manifest["cases"]))


if __name__ == "__main__":
    main()

"""Generates tests/golden/*.arrow + manifest.json by RUNNING THE REFERENCE.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

Every case feeds seeded synthetic blocks through the reference's own
``SortTaskSpec`` / ``ShuffleTaskSpec`` / ``SortAggregateTaskSpec`` ``map`` and
``reduce`` staticmethods (python/ray/data/_internal/planner/exchange/*.py),
scheduled as the pull-based scheduler does (all maps, then reducer j gets
column j of the map outputs -- pull_based_shuffle_task_scheduler.py:84-120), and
stores inputs and outputs as Arrow IPC files.  The fixtures travel to the GPU
box; the reference does not.
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


def _write(name: str, table: pa.Table) -> str:
    path = os.path.join(HERE, name)
    with pa.OSFile(path, "wb") as f:
        with pa.ipc.new_file(f, table.schema) as w:
            w.write_table(table)
    return name


def _payload(rng, n, width=None):
    if width is None:
        lens = rng.integers(0, 40, n)
    else:
        lens = np.full(n, width)
    return pa.array([rng.bytes(int(l)) for l in lens], type=pa.binary())


def make_blocks(rng, sizes, kind):
    blocks = []
    for n in sizes:
        if kind == "graysort":
            # int64 key + 90-byte binary payload (release/nightly_tests/dataset/
            # sort_benchmark.py:39-48 shape, 90 B as BASELINE.json config 2)
            t = pa.table({
                "c_0": pa.array(rng.integers(0, 2**63 - 1, n, dtype=np.int64)),
                "c_1": _payload(rng, n, 90),
            })
        elif kind == "dups":
            t = pa.table({
                "k": pa.array(rng.integers(-20, 20, n, dtype=np.int64)),
                "w": pa.array(rng.integers(0, 5, n, dtype=np.int32)),
                "f": pa.array(rng.standard_normal(n)),
                "p": _payload(rng, n),
            })
        elif kind == "float_nan_null":
            f = rng.standard_normal(n)
            f[rng.random(n) < 0.1] = np.nan
            f[rng.random(n) < 0.05] = np.inf
            f[rng.random(n) < 0.05] = -0.0
            t = pa.table({
                "f": pa.array(f, mask=rng.random(n) < 0.1),
                "i": pa.array(rng.integers(-3, 3, n, dtype=np.int64), mask=rng.random(n) < 0.2),
                "row": pa.array(np.arange(n, dtype=np.int32)),
            })
        elif kind == "agg":
            t = pa.table({
                "k": pa.array(rng.integers(0, 37, n, dtype=np.int32)),
                "v": pa.array(rng.standard_normal(n).astype(np.float32), mask=rng.random(n) < 0.1),
                "w": pa.array(rng.integers(-100, 100, n, dtype=np.int64)),
            })
        else:
            raise ValueError(kind)
        blocks.append(t)
    return blocks


def main():
    R = ref_shim.load()
    manifest = {"generator": "tests/golden/make_golden.py", "cases": []}

    def sort_case(name, kind, sizes, columns, descending, boundaries, seed):
        rng = np.random.default_rng(seed)
        blocks = make_blocks(rng, sizes, kind)
        sk = R.SortKey(key=columns, descending=descending)
        nred = len(boundaries) + 1
        maps = [R.SortTaskSpec.map(i, b, nred, boundaries, sk) for i, b in enumerate(blocks)]
        outs = [R.SortTaskSpec.reduce(sk, "default", *[m[j] for m in maps])[0] for j in range(nred)]
        case = {
            "name": name, "op": "sort", "columns": columns, "descending": descending,
            "boundaries": [list(b) for b in boundaries],
            "inputs": [_write(f"{name}.in{i}.arrow", b) for i, b in enumerate(blocks)],
            "map_outputs": [[_write(f"{name}.map{i}_{j}.arrow", m[j]) for j in range(nred)]
                            for i, m in enumerate(maps)],
            "outputs": [_write(f"{name}.out{j}.arrow", o) for j, o in enumerate(outs)],
        }
        manifest["cases"].append(case)

    sort_case("sort_gray_asc", "graysort", [700, 0, 513, 300], ["c_0"], [False],
              [(2**61,), (2**62,), (3 * 2**61,)], 11)
    sort_case("sort_gray_desc", "graysort", [700, 0, 513, 300], ["c_0"], [True],
              [(3 * 2**61,), (2**62,), (2**61,)], 12)
    sort_case("sort_dups_asc", "dups", [400, 300, 1], ["k"], [False], [(-10,), (0,), (0,), (10,)], 13)
    sort_case("sort_dups_desc", "dups", [400, 300, 1], ["k"], [True], [(10,), (0,), (-10,)], 14)
    sort_case("sort_multi_mixed", "dups", [400, 300, 250], ["k", "w"], [False, True],
              [(-10, 3), (0, 2), (0, 0), (10, 4)], 15)
    sort_case("sort_multi_desc_first", "dups", [400, 300, 250], ["k", "w"], [True, False],
              [(10, 0), (0, 2), (-10, 4)], 16)
    sort_case("sort_float_nan_null_asc", "float_nan_null", [300, 200], ["f"], [False],
              [(-0.5,), (0.0,), (0.5,)], 17)
    sort_case("sort_float_nan_null_desc_nobounds", "float_nan_null", [300, 200], ["f"], [True], [], 18)
    sort_case("sort_int_null_two_keys", "float_nan_null", [300, 200], ["i", "f"], [False, False],
              [(-1, 0.0), (1, 0.0)], 19)

    def shuffle_case(name, sizes, nout, seed):
        rng = np.random.default_rng(seed)
        blocks = make_blocks(rng, sizes, "dups")
        maps = [R.ShuffleTaskSpec.map(i, b, nout, 1 << 30, None, True, seed)
                for i, b in enumerate(blocks)]
        outs = [R.ShuffleTaskSpec.reduce(True, seed, *[m[j] for m in maps])[0] for j in range(nout)]
        manifest["cases"].append({
            "name": name, "op": "shuffle", "seed": seed, "num_outputs": nout,
            "inputs": [_write(f"{name}.in{i}.arrow", b) for i, b in enumerate(blocks)],
            "outputs": [_write(f"{name}.out{j}.arrow", o) for j, o in enumerate(outs)],
        })

    shuffle_case("shuffle_seed42", [257, 100, 3], 3, 42)

    def agg_case(name, sizes, boundaries, seed):
        rng = np.random.default_rng(seed)
        blocks = make_blocks(rng, sizes, "agg")
        A = R.aggregate
        aggs = [A.Count(), A.Sum("v"), A.Mean("v"), A.Min("v"), A.Max("v"), A.Std("v"), A.Sum("w")]
        spec = [["count"], ["sum", "v"], ["mean", "v"], ["min", "v"], ["max", "v"],
                ["std", "v", 1], ["sum", "w"]]
        sk = R.SortKey(key="k")
        nred = len(boundaries) + 1
        maps = [R.SortAggregateTaskSpec.map(i, b, nred, boundaries, sk, aggs)
                for i, b in enumerate(blocks)]
        outs = [R.SortAggregateTaskSpec.reduce(sk, aggs, "default", *[m[j] for m in maps])[0]
                for j in range(nred)]
        manifest["cases"].append({
            "name": name, "op": "aggregate", "key": "k", "aggs": spec,
            "boundaries": [list(b) for b in boundaries],
            "inputs": [_write(f"{name}.in{i}.arrow", b) for i, b in enumerate(blocks)],
            "outputs": [_write(f"{name}.out{j}.arrow", o) for j, o in enumerate(outs)],
        })

    agg_case("agg_int32key", [500, 300, 411], [(10,), (20,), (30,)], 21)

    with open(os.path.join(HERE, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    print("wrote", len(manifest["cases"]), "cases")


if __name__ == "__main__":
    main()

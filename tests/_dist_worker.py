"""Worker of tests/test_distributed.py: one rank of a world_size-N `gloo` run of the
multi-GPU sort on the host-memory test double (host logic + collectives, no GPU)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import pyarrow as pa
import torch.distributed as dist

from _fake_cabi import FakeLib
from _golden import tables_equal
from cluster_anywhere_b200 import SortKey, _cabi
from cluster_anywhere_b200.exchange import DeviceExchange
from oracle import exchange_oracle as O


def make_block(rng, n):
    k = rng.integers(-50, 50, n)
    f = rng.standard_normal(n)
    f[rng.random(n) < 0.1] = np.nan
    s = [None if rng.random() < 0.1 else ("s%d" % rng.integers(0, 1000)) * int(rng.integers(0, 4))
         for _ in range(n)]
    return pa.table({
        "k": pa.array(k, pa.int32()),
        "f": pa.array(f, pa.float64(), mask=rng.random(n) < 0.1),
        "s": pa.array(s, pa.string()),
        "b": pa.array(rng.random(n) < 0.5),
        "p": pa.array([bytes(rng.integers(0, 256, 10, dtype=np.uint8)) for _ in range(n)], pa.binary()),
        "q": pa.array([bytes(rng.integers(0, 256, 48, dtype=np.uint8)) for _ in range(n)], pa.binary()),
    })


def main():
    if "--gpu" in sys.argv:  # real kernels + NCCL + NVLink peer stores, one rank per GPU
        import torch

        lr = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(lr)
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    else:
        _cabi._install_test_double(FakeLib())
        dist.init_process_group("gloo")
    r, w = dist.get_rank(), dist.get_world_size()
    rng = np.random.default_rng(7)  # same stream on every rank: all ranks know all blocks
    sizes = (1000, 0, 300, 2000, 17) if w == 2 else (400, 0, 300, 0, 0, 900, 17)
    all_blocks = [make_block(rng, n) for n in sizes]
    cut = [len(all_blocks) * g // w for g in range(w + 1)]
    if "--empty-rank" in sys.argv:
        cut = [0] + [len(all_blocks)] * w  # rank 0 brings everything, the others nothing
    mine = all_blocks[cut[r]:cut[r + 1]]
    ex = DeviceExchange()
    full = pa.concat_tables([b for b in all_blocks if b.num_rows])
    cases = [(["k"], [False], None), (["k", "f"], [True, False], None), (["f"], [True], None),
             (["f", "k"], [True, True], None), (["s"], [False], None), (["s"], [True], None),
             (["s"], [False], [("s1",), ("s5",), (None,)]),
             (["k"], [False], [(-20,), (0,), (0,), (31,)]), (["k"], [True], [(10,), (-10,)]),
             # key tuples that mix numeric and string columns (xchg_*_mixed)
             (["k", "s"], [False, False], None), (["s", "k"], [False, True], None),
             (["k", "s"], [False, False], [(-20, "s1"), (0, None), (10, "s5"), (10, "s5s5")])]
    for cols, desc, bounds in cases:
        sk = SortKey(cols, desc)
        out = ex.sort_blocks_distributed(mine, sk, boundaries=bounds, seed=3)
        outs = [None] * w
        dist.all_gather_object(outs, out)
        got = [b for part in outs for b in part]
        exp = O.sort_table(full, cols, desc)
        nonempty = [g for g in got if g.num_rows]
        assert tables_equal(pa.concat_tables(nonempty), exp), (cols, desc, bounds)
        if bounds is None:
            assert len(got) == len(all_blocks), len(got)
        else:
            # per-block contents = the reference's range partitions of the sorted rows
            parts = O.find_partitions(exp, bounds, cols, desc)
            assert len(got) == len(parts)
            for g, e in zip(got, parts):
                assert g.num_rows == e.num_rows
                if e.num_rows:
                    assert tables_equal(g, e)
    # the chunked fast path (one fixed-width key without nulls, fixed / uniform payload): taken on
    # real GPUs (peer memory); under gloo the same input goes through the all-to-all path
    rng2 = np.random.default_rng(21)

    def plain_block(n):
        pay = rng2.integers(0, 256, n * 90, dtype=np.uint8)
        offs = np.arange(n + 1, dtype=np.int32) * 90
        return pa.table({"k": pa.array(rng2.integers(-10 ** 9, 10 ** 9, n), pa.int64()),
                         "v": pa.array(rng2.integers(0, 100, n), pa.int32()),
                         "p": pa.Array.from_buffers(pa.binary(), n, [None, pa.py_buffer(offs), pa.py_buffer(pay)])})

    plain = [plain_block(n) for n in ((30000, 5, 20000, 41000, 9000, 100) if w == 2 else (3000, 5, 2000, 4100, 900, 100, 77))]
    pcut = [len(plain) * g // w for g in range(w + 1)]
    pmine = plain[pcut[r]:pcut[r + 1]]
    pfull = pa.concat_tables(plain)
    for desc, bounds in ((False, None), (True, None), (False, [(-5 * 10 ** 8,), (0,), (123456789,)])):
        before = getattr(ex, "chunked_steps", 0)
        out = ex.sort_blocks_distributed(pmine, SortKey("k", desc), boundaries=bounds, seed=5)
        if "--gpu" in sys.argv:
            assert getattr(ex, "chunked_steps", 0) == before + 1, "chunked path not taken on the GPU"
        outs = [None] * w
        dist.all_gather_object(outs, out)
        got = [b for part in outs for b in part]
        exp = O.sort_table(pfull, ["k"], [desc])
        assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), exp), ("plain", desc, bounds)
        if bounds is not None:
            parts = O.find_partitions(exp, bounds, ["k"], [desc])
            assert [g.num_rows for g in got] == [e.num_rows for e in parts]
    # random_shuffle: the G-rank result is the single-process result, take(P^-1(0..N))
    for seed, nout in ((42, None), (7, 3)):
        out = ex.random_shuffle_blocks(mine, seed, num_outputs=nout)
        outs = [None] * w
        dist.all_gather_object(outs, out)
        got = [b for part in outs for b in part]
        n = full.num_rows
        idx = FakeLib()._feistel(seed, n, np.arange(n, dtype=np.uint64), True).astype(np.int64)
        assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]), full.take(idx)), seed
        r_total = len(all_blocks) if nout is None else nout
        assert [g.num_rows for g in got] == [(j + 1) * n // r_total - j * n // r_total
                                             for j in range(r_total)]
    # groupby().aggregate(): local hash aggregate, partials over the all-to-all, merge
    from cluster_anywhere_b200 import Count, Max, Mean, Min, Std, Sum
    from test_aggregate import assert_tables_close

    aggs = [Count(), Sum("f"), Mean("f"), Min("k"), Max("f"), Std("f", ddof=1)]
    exp = O.groupby_aggregate(all_blocks, "k", [("count",), ("sum", "f"), ("mean", "f"), ("min", "k"),
                                                 ("max", "f"), ("std", "f", 1)])
    for bounds in (None, [(-20,), (0,), (31,)]):
        out = ex.aggregate_blocks(mine, "k", aggs, boundaries=bounds, seed=1)
        outs = [None] * w
        dist.all_gather_object(outs, out)
        got = [b for part in outs for b in part]
        assert len(got) == (len(all_blocks) if bounds is None else len(bounds) + 1)
        assert_tables_close(pa.concat_tables([g for g in got if g.num_rows]), exp, exact=("k",))
        if bounds is not None:
            ks = [g["k"].to_pylist() for g in got]
            assert all(k is None or k < -20 for k in ks[0]) and all(-20 <= k < 0 for k in ks[1])
    # general keys (string key; int + string): sorted grouping, partials with string key
    # columns over the exchange, mixed-key splitters
    for gkey in ("s", ["k", "s"]):
        aggs2 = [Count(), Sum("f"), Mean("f")]
        exp2 = O.groupby_aggregate(all_blocks, gkey, [("count",), ("sum", "f"), ("mean", "f")])
        out = ex.aggregate_blocks(mine, gkey, aggs2, seed=2)
        outs = [None] * w
        dist.all_gather_object(outs, out)
        got = [b for part in outs for b in part]
        assert len(got) == len(all_blocks)
        keys = tuple([gkey] if isinstance(gkey, str) else gkey)
        assert_tables_close(pa.concat_tables([g for g in got if g.num_rows], promote_options="permissive"),
                            exp2, exact=keys)
    # the planner's device strategy over the process group: this rank's bundles in, its run
    # of the sorted output blocks out
    from cluster_anywhere_b200 import planner as P

    pctx = P.DataContext(shuffle_strategy=P.ShuffleStrategy.B200_DEVICE_EXCHANGE)
    bundles, stats = P.execute_all_to_all(P.Sort(P.InputData(), SortKey(["k", "f"], [False, True])),
                                          [P.RefBundle.from_block(b) for b in mine], pctx)
    outs = [None] * w
    dist.all_gather_object(outs, [b for rb in bundles for b in rb.block_refs])
    got = [b for part in outs for b in part]
    assert len(got) == len(all_blocks) and set(stats) == {"map", "reduce"}
    assert tables_equal(pa.concat_tables([g for g in got if g.num_rows]),
                        O.sort_table(full, ["k", "f"], [False, True]))
    bundles, _ = P.execute_all_to_all(P.Aggregate(P.InputData(), "k", [Count(), Sum("f")]),
                                      [P.RefBundle.from_block(b) for b in mine], pctx)
    outs = [None] * w
    dist.all_gather_object(outs, [b for rb in bundles for b in rb.block_refs])
    got = pa.concat_tables([b for part in outs for b in part if b.num_rows], promote_options="permissive")
    assert got["count()"].to_pylist() == O.groupby_aggregate(all_blocks, "k", [("count",)])["count()"].to_pylist()
    bundles, _ = P.execute_all_to_all(P.RandomShuffle(P.InputData(), seed=11),
                                      [P.RefBundle.from_block(b) for b in mine], pctx)
    outs = [None] * w
    dist.all_gather_object(outs, [b for rb in bundles for b in rb.block_refs])
    got = pa.concat_tables([b for part in outs for b in part if b.num_rows])
    assert got.num_rows == full.num_rows and sorted(got["k"].to_pylist()) == sorted(full["k"].to_pylist())
    out = ex.aggregate_blocks(mine, None, [Count(), Sum("k")])
    outs = [None] * w
    dist.all_gather_object(outs, out)
    got = [b for part in outs for b in part]
    assert len(got) == 1 and got[0]["count()"].to_pylist() == [full.num_rows]
    dist.barrier()
    if r == 0:
        print("DIST-OK", w)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

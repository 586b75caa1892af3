#!/usr/bin/env python
"""Secondary measurements: random_shuffle and groupby().aggregate() on the
per-GPU shards of BASELINE.json configs[2] / configs[3], inputs resident in HBM, CUDA
events, per-kernel device time from the library's profiling hooks.  One JSON line per op.

  python scripts/bench_ops.py [--steps K] [--warmup W]
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

KNAMES = {0: "k_normalize_hist", 1: "k_scan_hist", 2: "k_onesweep", 3: "k_gather_rows", 4: "k_gather_fixed",
          5: "k_gather_bitmap", 6: "k_lengths_scan", 7: "k_partition", 8: "k_offsets_*", 9: "misc",
          10: "k_merge", 11: "k_shuffle", 12: "k_aggregate", 13: "k_local_sort", 14: "k_msd_aux"}


def kernels(lib, steps):
    from cluster_anywhere_b200 import _cabi

    out = []
    for kid, name in KNAMES.items():
        tot, cnt = ctypes.c_double(), ctypes.c_uint64()
        _cabi.check(lib.xchg_profile_read(kid, ctypes.byref(tot), ctypes.byref(cnt)))
        if cnt.value:
            out.append({"kernel": name, "launches_per_step": cnt.value / steps, "ms_per_step": tot.value / steps})
    return sorted(out, key=lambda k: -k["ms_per_step"])


def timed(fn, steps, warmup, lib):
    import torch

    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    lib.xchg_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        out = fn(100 + i)
        del out
    e1.record()
    torch.cuda.synchronize()
    ks = kernels(lib, steps)
    lib.xchg_profile_enable(0)
    return e0.elapsed_time(e1) / steps, ks


def range_sort(args, rt, ex, lib, peak):
    """configs[0]: ray.data.range(10_000_000).sort('id', descending=True): one int64 column in 76
    blocks (what _autodetect_parallelism gives on an 8-CPU host), ascending input and a
    shuffled variant.  80 MB: launch- and latency-bound on a B200; reported, not tuned for."""
    import pyarrow as pa
    import torch

    from cluster_anywhere_b200 import SortKey
    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceDataset

    n, blocks = 10_000_000, 76
    for name, ids in (("ascending input", torch.arange(n, dtype=torch.int64, device=rt.device)),
                      ("shuffled input (seed 1)", torch.randperm(n, device=rt.device,
                                                                 generator=torch.Generator(device=rt.device).manual_seed(1)))):
        col = dev.DeviceColumn(pa.int64(), n, "fixed", 8, ids.view(torch.uint8))
        table = dev.DeviceTable(["id"], [col], n, pa.schema([("id", pa.int64())]))
        ds = DeviceDataset(table, [(i * n // blocks, (i + 1) * n // blocks) for i in range(blocks)])
        sk = SortKey("id", descending=True)
        ms, ks = timed(lambda i: ex.sort(ds, sk, seed=i), args.steps, args.warmup, lib)
        gbs = 2 * n * 8 / ms / 1e6
        yield {"op": "sort(range)", "workload": f"range({n}).sort('id', descending=True), {blocks} blocks, "
               f"{name}: BASELINE configs[0]", "ms_per_step": ms, "rows_per_s": n / ms * 1e3,
               "gb_per_s_algorithmic": gbs, "frac_of_measured_hbm": gbs / peak, "kernels": ks}


def string_sort(args, rt, ex, lib, peak):
    """configs[4] at a stated sub-scale: Dataset.sort on a string key drawn Zipf(1.2) from a
    vocabulary of 8..31-byte strings, with a 16..256-byte binary payload."""
    import numpy as np
    import pyarrow as pa
    import torch

    from cluster_anywhere_b200 import SortKey
    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceDataset

    n, V = args.string_rows, args.string_vocab
    rng = np.random.default_rng(5)
    gen = torch.Generator(device=rt.device)
    gen.manual_seed(5)
    vlen = torch.randint(8, 32, (V,), device=rt.device, generator=gen, dtype=torch.int64)
    voff = torch.zeros(V + 1, dtype=torch.int64, device=rt.device)
    voff[1:] = torch.cumsum(vlen, 0)
    vbytes = int(voff[-1].item())
    vdata = torch.randint(97, 123, (vbytes + 64,), device=rt.device, generator=gen, dtype=torch.uint8)[:vbytes]
    vocab = dev.DeviceColumn(pa.string(), V, "var", 1, vdata, voff, None, 0)
    z = np.minimum(rng.zipf(1.2, n), V) - 1
    idx = torch.from_numpy(z.astype(np.int32)).to(rt.device)
    keycol = dev.take_column(rt, vocab, idx, n)
    plen = torch.randint(16, 257, (n,), device=rt.device, generator=gen, dtype=torch.int64)
    poff = torch.zeros(n + 1, dtype=torch.int64, device=rt.device)
    poff[1:] = torch.cumsum(plen, 0)
    pbytes = int(poff[-1].item())
    pdata = torch.empty(pbytes + 64, dtype=torch.uint8, device=rt.device)[:pbytes]
    pdata[: pbytes // 8 * 8].view(torch.int64)[:] = torch.arange(pbytes // 8, device=rt.device, dtype=torch.int64)
    del plen, idx
    paycol = dev.DeviceColumn(pa.large_binary(), n, "var", 1, pdata, poff, None, 0)
    table = dev.DeviceTable(["key", "payload"], [keycol, paycol], n,
                            pa.schema([("key", pa.string()), ("payload", pa.large_binary())]))
    blocks = 16
    ds = DeviceDataset(table, [(i * n // blocks, (i + 1) * n // blocks) for i in range(blocks)])
    sk = SortKey("key")
    kbytes = int(keycol.data.numel())
    ms, ks = timed(lambda i: ex.sort(ds, sk, seed=i), args.steps, args.warmup, lib)
    gbs = 2 * (kbytes + pbytes + 16 * (n + 1)) / ms / 1e6
    res = {"op": "sort(string key)", "workload": f"{n} rows, string key Zipf(1.2) over {V} distinct strings of "
           f"8-31 bytes ({kbytes / 1e9:.2f} GB), binary payload 16-256 bytes ({pbytes / 1e9:.2f} GB), {blocks} "
           "blocks: BASELINE configs[4] (500 M rows, 10 M distinct) at a sub-scale that fits one run's time",
           "ms_per_step": ms, "rows_per_s": n / ms * 1e3, "gb_per_s_algorithmic": gbs,
           "frac_of_measured_hbm": gbs / peak, "kernels": ks}
    del ds, table, keycol, paycol, pdata, poff, vocab
    torch.cuda.empty_cache()
    return res


def shuffle_op(args, rt, ex, lib, peak):
    """configs[2] shard: random_shuffle(seed=42) of 1 KiB tensor rows."""
    import pyarrow as pa
    import torch

    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceDataset

    n, w = args.shuffle_rows, 1024
    data = torch.empty(n * w + 64, dtype=torch.uint8, device=rt.device)[: n * w]
    data.view(torch.int64)[:] = torch.arange(n * w // 8, device=rt.device, dtype=torch.int64)  # row-id derived
    typ = pa.large_list(pa.uint8())
    col = dev.DeviceColumn(typ, n, "var", 1, data, None, None, 0, uniform=w)
    table = dev.DeviceTable(["data"], [col], n, pa.schema([("data", typ)]))
    blocks = 24
    ds = DeviceDataset(table, [(i * n // blocks, (i + 1) * n // blocks) for i in range(blocks)])
    ms, ks = timed(lambda i: ex.random_shuffle(ds, 42), args.steps, args.warmup, lib)
    gbs = 2 * n * w / ms / 1e6
    res = {"op": "random_shuffle", "workload": f"{n} rows x 1 KiB tensor rows ({n * w / 1e9:.1f} GB), seed 42, "
           f"{blocks} blocks: the per-GPU shard of BASELINE configs[2] (200 M rows at 8 GPUs)",
           "ms_per_step": ms, "rows_per_s": n / ms * 1e3, "gb_per_s_algorithmic": gbs,
           "frac_of_measured_hbm": gbs / peak, "kernels": ks}
    del ds, table, col, data
    torch.cuda.empty_cache()
    return res


def aggregate_op(args, rt, ex, lib, peak):
    """configs[3] shard: groupby(int32 key).aggregate(Sum, Mean, Count) over float32 rows."""
    import pyarrow as pa
    import torch

    from cluster_anywhere_b200 import Count, Mean, Sum
    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceDataset

    n, g = args.agg_rows, args.agg_groups
    gen = torch.Generator(device=rt.device)
    gen.manual_seed(4)
    k = torch.randint(0, g, (n,), dtype=torch.int32, device=rt.device, generator=gen)
    v = torch.randn(n, dtype=torch.float32, device=rt.device, generator=gen)
    kc = dev.DeviceColumn(pa.int32(), n, "fixed", 4, k.view(torch.uint8))
    vc = dev.DeviceColumn(pa.float32(), n, "fixed", 4, v.view(torch.uint8))
    table = dev.DeviceTable(["k", "v"], [kc, vc], n, pa.schema([("k", pa.int32()), ("v", pa.float32())]))
    blocks = 16
    ds = DeviceDataset(table, [(i * n // blocks, (i + 1) * n // blocks) for i in range(blocks)])
    aggs = [Sum("v"), Mean("v"), Count()]
    ms, ks = timed(lambda i: ex.aggregate(ds, "k", aggs, seed=i), args.steps, args.warmup, lib)
    out = ex.aggregate(ds, "k", aggs)
    groups = out.num_rows
    gbs = (8 * n + 32 * groups) / ms / 1e6
    res = {"op": "groupby.aggregate", "workload": f"{n} rows (int32 key, float32 value), {groups} groups, "
           f"Sum/Mean/Count, {blocks} blocks: the per-GPU shard of BASELINE configs[3] (2 B rows at 8 GPUs)",
           "ms_per_step": ms, "rows_per_s": n / ms * 1e3, "gb_per_s_algorithmic": gbs,
           "frac_of_measured_hbm": gbs / peak, "kernels": ks}
    del ds, table, kc, vc, k, v, out
    torch.cuda.empty_cache()
    return res


class OpArgs:
    """Default sizes of the secondary measurements (bench.py's ``ops`` key uses these)."""

    def __init__(self, steps=5, warmup=3):
        self.steps, self.warmup = steps, warmup
        self.shuffle_rows = 25_000_000
        self.agg_rows = 250_000_000
        self.agg_groups = 10_000_000
        self.string_rows = 40_000_000
        self.string_vocab = 800_000


def run_ops(rt, ex, lib, peak, steps=5, warmup=3, only=()):
    """The secondary operators on one GPU -> list of result dicts."""
    args = OpArgs(steps, warmup)
    out = []
    if not only or "shuffle" in only:
        out.append(shuffle_op(args, rt, ex, lib, peak))
    if not only or "aggregate" in only:
        out.append(aggregate_op(args, rt, ex, lib, peak))
    if not only or "strings" in only:
        out.append(string_sort(args, rt, ex, lib, peak))
    if not only or "c1" in only:
        out.extend(range_sort(args, rt, ex, lib, peak))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--shuffle-rows", type=int, default=25_000_000)
    ap.add_argument("--agg-rows", type=int, default=250_000_000)
    ap.add_argument("--agg-groups", type=int, default=10_000_000)
    ap.add_argument("--string-rows", type=int, default=40_000_000)
    ap.add_argument("--string-vocab", type=int, default=800_000)
    ap.add_argument("--only", default="", help="comma list of ops to run: c1,shuffle,aggregate,strings")
    args = ap.parse_args()
    import __graft_entry__ as ge

    ge.build()
    from cluster_anywhere_b200 import _cabi
    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceExchange

    lib = _cabi.load()
    rt = dev.Runtime()
    ex = DeviceExchange(rt)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    only = set(x for x in args.only.split(",") if x)
    if not only or "c1" in only:
        for r in range_sort(args, rt, ex, lib, peak):
            print(json.dumps(r), flush=True)
    if not only or "strings" in only:
        print(json.dumps(string_sort(args, rt, ex, lib, peak)), flush=True)
    if not only or "shuffle" in only:
        print(json.dumps(shuffle_op(args, rt, ex, lib, peak)), flush=True)
    if not only or "aggregate" in only:
        print(json.dumps(aggregate_op(args, rt, ex, lib, peak)), flush=True)


if __name__ == "__main__":
    main()

"""BASELINE.json configs[1] at its full per-GPU size (125 M GraySort rows: int64 key + 90-byte
payload, 12.25 GB), checked through size-independent properties instead of the oracle:

* key order: the output keys are non-decreasing, block after block;
* row integrity: every payload row still carries the key it was generated from (bytes 0-7)
  and the key-derived filler (so no row was torn, duplicated from another row, or lost);
* row multiset: the sum of the keys (mod 2^64) is that of the input, the row count is unchanged;
* partitioning: the block cut points are consistent with the keys around them;
* idempotence: sorting the sorted dataset changes nothing.
The same properties for the all-global LSD passes (xchg_set_sort_algorithm(1)), which must
produce the identical permutation."""
import numpy as np
import pyarrow as pa
import pytest
import torch

from cluster_anywhere_b200 import SortKey
from cluster_anywhere_b200 import device as dev
from cluster_anywhere_b200.exchange import DeviceDataset, DeviceExchange

pytestmark = pytest.mark.gpu

N = 125_000_000
W = 90


def _make(rt, n):
    g = torch.Generator(device=rt.device)
    g.manual_seed(2)
    keys = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=rt.device, generator=g)
    data = torch.empty(n * W + 64, dtype=torch.uint8, device=rt.device)[: n * W]
    rows = data.view(n, W)
    kb = keys.view(torch.uint8).view(n, 8)
    rows[:, 0:8] = kb
    for j in range(8, W, 8):  # filler: the key bytes rotated by the column index
        w = min(8, W - j)
        rows[:, j:j + w] = torch.roll(kb, shifts=j // 8, dims=1)[:, :w]
    kc = dev.DeviceColumn(pa.int64(), n, "fixed", 8, keys.view(torch.uint8))
    pc = dev.DeviceColumn(pa.binary(), n, "var", 1, data, None, None, 0, uniform=W)
    table = dev.DeviceTable(["c_0", "c_1"], [kc, pc], n, pa.schema([("c_0", pa.int64()), ("c_1", pa.binary())]))
    blocks = 12
    return keys, DeviceDataset(table, [(i * n // blocks, (i + 1) * n // blocks) for i in range(blocks)])


def _check(rt, keys_in, out):
    n = keys_in.numel()
    assert out.num_rows == n and len(out.block_ranges) == 12
    ok = out.table.column("c_0").data.view(torch.int64)[:n]
    assert bool((ok[1:] >= ok[:-1]).all()), "keys out of order"
    assert int(ok.sum().item()) == int(keys_in.sum().item())
    rows = out.table.column("c_1").data[: n * W].view(n, W)
    kb = ok.view(torch.uint8).view(n, 8)
    assert bool((rows[:, 0:8] == kb).all()), "payload separated from its key"
    for j in (8, 40, 88):
        w = min(8, W - j)
        assert bool((rows[:, j:j + w] == torch.roll(kb, shifts=j // 8, dims=1)[:, :w]).all()), f"payload bytes {j}.."
    last = 0
    for lo, hi in out.block_ranges:
        assert lo == last and hi >= lo
        last = hi
    assert last == n
    return ok


def test_graysort_shard_full_size_properties():
    rt = dev.default_runtime()
    keys, ds = _make(rt, N)
    ex = DeviceExchange(rt)
    sk = SortKey("c_0")
    out = ex.sort(ds, sk, seed=1)
    ok = _check(rt, keys, out)
    # block sizes: sampled boundaries over uniform keys -> no block far from n / 12
    sizes = [hi - lo for lo, hi in out.block_ranges]
    assert max(sizes) < 3 * N // 12
    # the all-global LSD passes give the same rows in the same order
    rt.call("xchg_set_sort_algorithm", 1, launches=0)
    try:
        out2 = ex.sort(ds, sk, boundaries=[], seed=1)
    finally:
        rt.call("xchg_set_sort_algorithm", 0, launches=0)
    assert torch.equal(out2.table.column("c_0").data[: N * 8], out.table.column("c_0").data[: N * 8])
    assert torch.equal(out2.table.column("c_1").data[: N * W], out.table.column("c_1").data[: N * W])
    del out2
    # idempotence: sorting the sorted rows changes nothing
    again = ex.sort(DeviceDataset(out.table, out.block_ranges), sk, boundaries=[], seed=1)
    assert torch.equal(again.table.column("c_1").data[: N * W], out.table.column("c_1").data[: N * W])
    assert torch.equal(again.table.column("c_0").data.view(torch.int64)[:N], ok)


def test_random_shuffle_shard_full_size_properties():
    """configs[2] per-GPU shard: 25 M rows of 1 KiB.  The output is a permutation of the input
    rows (every row carries its id), rows stay intact, the same seed gives the same output,
    another seed a different one, and the rows of an input block do not stay together."""
    rt = dev.default_runtime()
    n, w = 25_000_000, 1024
    data = torch.empty(n * w + 64, dtype=torch.uint8, device=rt.device)[: n * w]
    words = data.view(torch.int64).view(n, w // 8)
    ids = torch.arange(n, dtype=torch.int64, device=rt.device)
    words[:] = ids[:, None]                      # every 8-byte word of row i holds i ...
    words[:, 1] = ids * 0x9E3779B97F4A7C15 % (2**62)  # ... but word 1: a hash of i
    typ = pa.large_list(pa.uint8())
    col = dev.DeviceColumn(typ, n, "var", 1, data, None, None, 0, uniform=w)
    table = dev.DeviceTable(["data"], [col], n, pa.schema([("data", typ)]))
    blocks = 24
    ds = DeviceDataset(table, [(i * n // blocks, (i + 1) * n // blocks) for i in range(blocks)])
    ex = DeviceExchange(rt)
    out = ex.random_shuffle(ds, 42)
    assert out.num_rows == n and len(out.block_ranges) == blocks
    ow = out.table.column("data").data[: n * w].view(torch.int64).view(n, w // 8)
    got = ow[:, 0].clone()
    assert bool((ow[:, 5] == got).all()) and bool((ow[:, 127] == got).all()), "row torn"
    assert bool((ow[:, 1] == got * 0x9E3779B97F4A7C15 % (2**62)).all()), "row torn"
    assert bool((torch.sort(got).values == ids).all()), "not a permutation of the input rows"
    # rows of input block 0 are spread over the output, not contiguous
    first_block = (got < n // blocks)
    per_out_block = torch.stack([first_block[lo:hi].sum() for lo, hi in out.block_ranges]).cpu().numpy()
    assert per_out_block.min() > 0 and per_out_block.max() < 2 * (n // blocks) // blocks
    del out, ow
    again = ex.random_shuffle(ds, 42).table.column("data").data[: n * w].view(torch.int64).view(n, w // 8)[:, 0]
    assert torch.equal(again, got), "same seed, different output"
    del again
    other = ex.random_shuffle(ds, 43).table.column("data").data[: n * w].view(torch.int64).view(n, w // 8)[:, 0]
    assert not torch.equal(other, got)


def test_groupby_aggregate_shard_full_size_properties():
    """configs[3] per-GPU shard: 250 M rows, int32 key with 10 M distinct values, float32 value,
    Sum / Mean / Count.  Counts add up to the row count and equal a torch.bincount of the keys,
    the group keys are the sorted distinct keys, sums match an fp64 index_add within 1e-6
    relative (the stated tolerance), mean = sum / count."""
    from cluster_anywhere_b200 import Count, Mean, Sum

    rt = dev.default_runtime()
    n, g = 250_000_000, 10_000_000
    gen = torch.Generator(device=rt.device)
    gen.manual_seed(4)
    k = torch.randint(0, g, (n,), dtype=torch.int32, device=rt.device, generator=gen)
    v = torch.randn(n, dtype=torch.float32, device=rt.device, generator=gen)
    kc = dev.DeviceColumn(pa.int32(), n, "fixed", 4, k.view(torch.uint8))
    vc = dev.DeviceColumn(pa.float32(), n, "fixed", 4, v.view(torch.uint8))
    table = dev.DeviceTable(["k", "v"], [kc, vc], n, pa.schema([("k", pa.int32()), ("v", pa.float32())]))
    ds = DeviceDataset(table, [(i * n // 16, (i + 1) * n // 16) for i in range(16)])
    out = DeviceExchange(rt).aggregate(ds, "k", [Sum("v"), Mean("v"), Count()], seed=1)
    G = out.num_rows
    t = out.table
    keys = t.column("k").data.view(torch.int64)[:G]
    cnt = t.column("count()").data.view(torch.int64)[:G]
    sm = t.column("sum(v)").data.view(torch.float64)[:G]
    mean = t.column("mean(v)").data.view(torch.float64)[:G]
    exp_cnt = torch.bincount(k.to(torch.int64), minlength=g)
    present = torch.nonzero(exp_cnt).flatten()
    assert G == present.numel() and torch.equal(keys, present)
    assert torch.equal(cnt, exp_cnt[present]) and int(cnt.sum().item()) == n
    exp_sum = torch.zeros(g, dtype=torch.float64, device=rt.device).index_add_(0, k.to(torch.int64), v.to(torch.float64))
    es = exp_sum[present]
    tol = 1e-6 * torch.maximum(es.abs(), sm.abs()) + 1e-9
    assert bool(((sm - es).abs() <= tol).all())
    assert bool(((mean - sm / cnt.to(torch.float64)).abs() <= 1e-12 * mean.abs() + 1e-15).all())
    last = 0
    for lo, hi in out.block_ranges:
        assert lo == last
        last = hi
    assert last == G and len(out.block_ranges) == 16

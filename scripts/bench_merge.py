#!/usr/bin/env python
"""A/B of the reduce side on one GPU: K6 (stable k-way merge of G sorted runs of radix keys,
xchg_merge_sorted_runs) against what the reference does and round 1 copied -- sorting the
concatenated runs again (xchg_sort_column).  G = 2, 4, 8 runs of N / G random 64-bit keys.
One JSON line per G; CUDA events, 5 repetitions after 2 warm-ups.

Reference: ArrowBlockAccessor.merge_sorted_blocks -> concat_and_sort
(python/ray/data/_internal/arrow_block.py:529-541, _internal/arrow_ops/transform_pyarrow.py:540-547)."""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=125_000_000)
    args = ap.parse_args()
    import torch

    import __graft_entry__ as ge

    ge.build()
    from cluster_anywhere_b200 import _cabi
    from cluster_anywhere_b200 import device as dev

    lib = _cabi.load()
    rt = dev.Runtime()
    n = args.rows

    def vp(t):
        return ctypes.c_void_p(t.data_ptr())

    def timed(fn, reps=5):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    g = torch.Generator(device=rt.device)
    g.manual_seed(1)
    for G in (2, 4, 8):
        per = n // G
        m = per * G
        keys = torch.randint(0, 2 ** 63 - 1, (m,), dtype=torch.int64, device=rt.device, generator=g)
        runs = torch.sort(keys.view(G, per), dim=1).values.reshape(-1).contiguous()  # G sorted runs back to back
        del keys
        starts = (ctypes.c_uint64 * (G + 1))(*[i * per for i in range(G + 1)])
        wsb = ctypes.c_uint64()
        _cabi.check(lib.xchg_merge_workspace_bytes(m, ctypes.byref(wsb)))
        ws = rt.empty(wsb.value)
        ko, vo = rt.empty(m, torch.int64), rt.empty(m, torch.int32)
        flag = rt.zeros(1, torch.int32)
        t_merge = timed(lambda: _cabi.check(lib.xchg_merge_sorted_runs(vp(runs), None, starts, G, vp(ko), vp(vo), vp(ws),
                                                                       wsb.value, vp(flag), rt.stream)))
        merged = ko.clone()
        sws = dev.sort_workspace_bytes(m)
        sw = rt.empty(sws)
        so, sn = rt.empty(m, torch.int32), rt.empty(m, torch.int64)
        t_sort = timed(lambda: _cabi.check(lib.xchg_sort_column(vp(runs), _cabi.U64, 0, None, 0, None, m, vp(so), vp(sn),
                                                                 vp(sw), sws, rt.stream)))
        same = bool(torch.equal(merged, sn)) and bool(torch.equal(vo, so)) and int(flag.item()) == 0
        rounds = (G - 1).bit_length()
        print(json.dumps({"runs": G, "keys": m, "merge_ms": round(t_merge, 3), "resort_ms": round(t_sort, 3),
                          "merge_rounds": rounds, "merge_gb_s_algorithmic": round(24 * m * rounds / t_merge / 1e6, 1),
                          "identical_output": same}), flush=True)
        del runs, ko, vo, so, sn, merged, ws, sw
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

"""Sort-only driver for profiling: n random 63-bit int64 keys, xchg_sort_column `reps`
times (CUDA-event time per call printed).  Used under ncu for the onesweep / local-sort
captures in profiles/."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cluster_anywhere_b200 import _cabi  # noqa: E402
from cluster_anywhere_b200 import device as dev  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=125_000_000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--bits", type=int, default=63)
    args = ap.parse_args()
    rt = dev.default_runtime()
    n = args.rows
    g = torch.Generator(device=rt.device)
    g.manual_seed(2)
    keys = torch.randint(0, min(2**args.bits, 2**63) - 1, (n,), dtype=torch.int64, device=rt.device, generator=g)
    ws_bytes = dev.sort_workspace_bytes(n)
    ws = rt.empty(ws_bytes)
    perm = rt.empty(n, torch.int32)
    norm = rt.empty(n, torch.int64)
    for r in range(args.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rt.call("xchg_sort_column", dev._vp(keys), _cabi.I64, 0, None, 0, None, n, dev._vp(perm), dev._vp(norm),
                dev._vp(ws), ws_bytes, rt.stream, launches=0)
        e1.record()
        torch.cuda.synchronize()
        print(f"rep {r}: {e0.elapsed_time(e1):.3f} ms", flush=True)
    import ctypes
    lib = _cabi.load()
    lib.xchg_profile_enable(1)
    rt.call("xchg_sort_column", dev._vp(keys), _cabi.I64, 0, None, 0, None, n, dev._vp(perm), dev._vp(norm),
            dev._vp(ws), ws_bytes, rt.stream, launches=0)
    torch.cuda.synchronize()
    names = {0: "normalize_hist", 1: "scan_hist", 2: "onesweep", 9: "misc", 13: "local_sort", 14: "msd_aux"}
    out = []
    for kid, name in names.items():
        tot, cnt = ctypes.c_double(), ctypes.c_uint64()
        lib.xchg_profile_read(kid, ctypes.byref(tot), ctypes.byref(cnt))
        if cnt.value:
            out.append(f"{name} x{cnt.value} {tot.value:.3f} ms")
    lib.xchg_profile_enable(0)
    print("  " + ", ".join(out), flush=True)
    s = norm[:: max(1, n // 1000)]
    assert bool((s[1:] >= s[:-1]).all())


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""What page-locking caller-owned wire bytes buys (cluster_anywhere_b200/wire.py): host->device
copy of a pageable buffer as it is, the cost of xchg_host_register on it, and the copy out of the
registered range.  One JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch

    import __graft_entry__ as ge

    ge.build()
    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200 import wire

    gb = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
    n = int(gb * (1 << 30))
    rt = dev.Runtime()
    host = np.ones(n, dtype=np.uint8)  # touched: every page resident
    d = torch.empty(n, dtype=torch.uint8, device=rt.device)
    src = torch.from_numpy(host)

    def copy_s():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        d.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    copy_s()
    pageable = min(copy_s() for _ in range(3))
    reg = wire.HostRegistration(rt)
    t0 = time.perf_counter()
    ok = reg.add(host)
    t_reg = time.perf_counter() - t0
    pinned_flag = bool(src.is_pinned())
    registered = min(copy_s() for _ in range(3))
    t0 = time.perf_counter()
    reg.release()
    t_unreg = time.perf_counter() - t0
    print(json.dumps({"bytes": n, "h2d_pageable_gb_s": round(n / pageable / 1e9, 1),
                      "register_s": round(t_reg, 3), "register_gb_s": round(n / t_reg / 1e9, 1),
                      "registered_ok": ok, "is_pinned_after_register": pinned_flag,
                      "h2d_registered_gb_s": round(n / registered / 1e9, 1), "unregister_s": round(t_unreg, 3)}))


if __name__ == "__main__":
    main()

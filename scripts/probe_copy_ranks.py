"""Host <-> device copy bandwidth per rank when R ranks copy at once (run under torchrun,
one rank per GPU; `--ranks 1,2,4,8` takes subsets of the launched ranks in turn).  Every
rank owns a 4 GiB pinned host buffer and a device buffer; a barrier, then all ACTIVE ranks
copy H2D (then D2H, then both at once) and the slowest active rank's time counts.  Shows
whether the end-to-end path of N ranks is bound by each GPU's PCIe link or by the host's
memory system (the N3 row: python/ray/_private/arrow_serialization.py:103 hands blocks over
in host memory)."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", default="1,2,4,8")
    ap.add_argument("--gib", type=int, default=4)
    args = ap.parse_args()
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    rank, world = dist.get_rank(), dist.get_world_size()
    n = args.gib << 30
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    t0 = time.perf_counter()
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    pin_s = time.perf_counter() - t0
    h.fill_(1)
    h2 = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    s2 = torch.cuda.Stream()
    pins = [None] * world
    dist.all_gather_object(pins, round(pin_s, 2))
    if rank == 0:
        print(json.dumps({"probe": f"cudaHostAlloc of {args.gib} GiB, seconds per rank, {world} ranks at once", "s": pins}),
              flush=True)

    def run(fn, active):
        best = 1e9
        for _ in range(3):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            if active:
                fn()
                torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            t = torch.tensor([dt if active else 0.0], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = min(best, float(t.item()))
        return best

    def both():
        with torch.cuda.stream(s2):
            h2.copy_(d, non_blocking=True)
        d.copy_(h, non_blocking=True)

    for r in [int(x) for x in args.ranks.split(",") if int(x) <= world]:
        active = rank < r
        res = {"ranks_copying": r, "gib_per_rank": args.gib}
        for name, fn, mult in (("h2d", lambda: d.copy_(h, non_blocking=True), 1),
                               ("d2h", lambda: h.copy_(d, non_blocking=True), 1), ("h2d+d2h", both, 2)):
            t = run(fn, active)
            res[name + "_gb_s_per_rank"] = round(mult * n / t / 1e9, 1)
            res[name + "_gb_s_total"] = round(mult * n * r / t / 1e9, 1)
        if rank == 0:
            print(json.dumps(res), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Probe: what does torch.distributed.all_to_all_single (NCCL grouped send/recv) reach on
this box for the sizes the sort exchange moves?  Run under torchrun with N ranks."""
import os
import time

import torch
import torch.distributed as dist


def main():
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    r, w = dist.get_rank(), dist.get_world_size()
    for total_gb in (1.0, 11.25):
        n = int(total_gb * 1e9) // w * w
        send = torch.empty(n, dtype=torch.uint8, device="cuda")
        recv = torch.empty(n, dtype=torch.uint8, device="cuda")
        splits = [n // w] * w
        for _ in range(3):
            dist.all_to_all_single(recv, send, splits, splits)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        iters = 5
        for _ in range(iters):
            dist.all_to_all_single(recv, send, splits, splits)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        off_gpu = n * (w - 1) / w
        if r == 0:
            print(f"[a2a] world={w} bytes/rank={n/1e9:.2f} GB: {ms:.2f} ms -> "
                  f"{off_gpu/ms/1e6:.1f} GB/s per direction leaving each GPU, {n/ms/1e6:.1f} GB/s incl. self part",
                  flush=True)
        # fresh allocation each time (what the exchange does today)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            rb = torch.empty(n + 4096 * _, dtype=torch.uint8, device="cuda")
            dist.all_to_all_single(rb[:n], send, splits, splits)
            del rb
        torch.cuda.synchronize()
        if r == 0:
            print(f"[a2a] with a fresh recv allocation per call: {(time.perf_counter()-t0)/3*1e3:.2f} ms", flush=True)
        del send, recv
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

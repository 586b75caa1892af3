"""NVLink probe for the exchange's payload path (run under torchrun, 2+ ranks, one per GPU):
how fast does ONE GPU's egress go when every rank moves ROWS (90-byte GraySort payloads,
random source rows) to / from its right-hand neighbour at the same time, by

  push_gather   K5 reading local rows at random, storing them densely into the peer (what the
                exchange does today), with the gather's CTA cap at 0 (= 24/SM), 1, 2, 3
  push_copy     contiguous 16-byte-per-thread stores into the peer (xchg_copy_segments)
  push_copy_tma the same through the TMA unit (cp.async.bulk global -> shared -> global)
  push_ce       cudaMemcpyAsync device-to-peer (copy engines)
  pull_gather   K5 reading the PEER's rows at random over NVLink, writing locally (the
                "key-first" alternative: every payload row moves once)
  mixed         K5 multi-destination: half the rows to the own arena, half to the peer
                (the 2-GPU exchange's real store pattern)

Prints one JSON line per variant on rank 0: GB/s of payload bytes per GPU (max time over
ranks), CUDA events around 3 repetitions after one warm-up.
"""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import __graft_entry__ as ge

    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    if lr != 0:
        import time

        time.sleep(0.5)
    ge.build()
    import torch.distributed._symmetric_memory as symm_mem

    from cluster_anywhere_b200 import _cabi
    from cluster_anywhere_b200 import device as dev

    rank, world = dist.get_rank(), dist.get_world_size()
    rt = dev.Runtime()
    W = 90
    n = int(os.environ.get("PROBE_ROWS", str(62_500_000)))  # 5.6 GB
    nbytes = n * W
    # symmetric: [table | landing zone]
    buf = symm_mem.empty(2 * nbytes + 4096, dtype=torch.uint8, device=rt.device)
    hdl = symm_mem.rendezvous(buf, group=dist.group.WORLD)
    ptrs = [int(p) for p in hdl.buffer_ptrs]
    peer = (rank + 1) % world
    table = buf[:nbytes]
    table.copy_(torch.randint(0, 256, (nbytes,), dtype=torch.uint8, device=rt.device))
    land_off = (nbytes + 255) // 256 * 256
    local_out = rt.empty(nbytes)
    g = torch.Generator(device=rt.device)
    g.manual_seed(rank)
    perm = torch.randperm(n, device=rt.device, generator=g).to(torch.int32)
    lib = _cabi.load()
    stream = rt.stream

    def vp(x):
        return ctypes.c_void_p(int(x))

    def timed(name, fn, nbytes_moved, reps=3):
        fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        t = torch.tensor([ms], device=rt.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        if rank == 0:
            print(json.dumps({"variant": name, "ms": round(float(t.item()), 3),
                              "gb_per_s_per_gpu": round(nbytes_moved / (float(t.item()) * 1e-3) / 1e9, 1),
                              "bytes": nbytes_moved, "world": world}), flush=True)

    def gather(in_ptr, out_ptr):
        _cabi.check(lib.xchg_gather_rows(vp(in_ptr), W, vp(perm.data_ptr()), n, vp(out_ptr), stream))

    if os.environ.get("PROBE_PAIR", "1" if world <= 2 else "0") == "1":
        for cap in (0, 1, 2, 3, 6):
            lib.xchg_set_gather_ctas_per_sm(cap)
            timed(f"push_gather cap={cap}", lambda: gather(table.data_ptr(), ptrs[peer] + land_off), nbytes)
            timed(f"pull_gather cap={cap}", lambda: gather(ptrs[peer], local_out.data_ptr()), nbytes)
        lib.xchg_set_gather_ctas_per_sm(0)
        timed("local_gather", lambda: gather(table.data_ptr(), local_out.data_ptr()), nbytes)

        def copy_seg(dst):
            src = (ctypes.c_void_p * 1)(table.data_ptr())
            d = (ctypes.c_void_p * 1)(dst)
            nb = (ctypes.c_uint64 * 1)(nbytes)
            _cabi.check(lib.xchg_copy_segments(1, src, d, nb, stream))

        timed("push_copy", lambda: copy_seg(ptrs[peer] + land_off), nbytes)
        timed("local_copy", lambda: copy_seg(local_out.data_ptr()), nbytes)
        # ctypes' getenv sees the C environment, not os.environ's copy: set it there
        libc = ctypes.CDLL(None)
        libc.setenv(b"XCHG_COPY_BULK", b"1", 1)
        timed("push_copy_tma", lambda: copy_seg(ptrs[peer] + land_off), nbytes)
        timed("local_copy_tma", lambda: copy_seg(local_out.data_ptr()), nbytes)
        libc.setenv(b"XCHG_COPY_BULK", b"0", 1)
        peer_view = hdl.get_buffer(peer, (nbytes,), torch.uint8, land_off)
        timed("push_ce", lambda: peer_view.copy_(table, non_blocking=True), nbytes)

        # mixed: half local, half remote in one multi-destination launch
        half = n // 2
        row0 = (ctypes.c_uint64 * 2)(0, half)
        rows = (ctypes.c_uint64 * 2)(half, n - half)
        outs = (ctypes.c_void_p * 2)(ptrs[rank] + land_off, ptrs[peer] + land_off + half * W)
        for cap in (0, 1, 2, 3):
            lib.xchg_set_gather_ctas_per_sm(cap)
            timed(f"mixed_multi cap={cap} (remote half counted)",
                  lambda: _cabi.check(lib.xchg_gather_rows_multi(vp(table.data_ptr()), W, vp(perm.data_ptr()), 2, row0, rows,
                                                                 outs, stream)), (n - half) * W)
        lib.xchg_set_gather_ctas_per_sm(0)

    # all-to-all: every rank stores an equal share of its rows into EVERY rank (itself
    # included) at the same time -- the exchange's real traffic pattern at world > 2
    if world > 2 or os.environ.get("PROBE_A2A", "0") == "1":
        share = n // world
        remote_bytes = share * W * (world - 1)
        order = [(rank + 1 + i) % world for i in range(world)]
        row0 = (ctypes.c_uint64 * world)(*[i * share for i in range(world)])
        rows = (ctypes.c_uint64 * world)(*[share] * world)
        outs = (ctypes.c_void_p * world)(*[ptrs[d] + land_off + rank * share * W for d in order])
        for cap in (0, 1, 2, 3, 6):
            lib.xchg_set_gather_ctas_per_sm(cap)
            timed(f"a2a_gather_multi cap={cap} (remote bytes counted)",
                  lambda: _cabi.check(lib.xchg_gather_rows_multi(vp(table.data_ptr()), W, vp(perm.data_ptr()), world,
                                                                 row0, rows, outs, stream)), remote_bytes)
        lib.xchg_set_gather_ctas_per_sm(0)

        def one_by_one():
            for i, d in enumerate(order):
                _cabi.check(lib.xchg_gather_rows(vp(table.data_ptr()), W, vp(perm.data_ptr() + 4 * i * share), share,
                                                 vp(ptrs[d] + land_off + rank * share * W), stream))

        timed("a2a_gather one launch per destination, rotated order (remote bytes counted)", one_by_one, remote_bytes)
        srcs = (ctypes.c_void_p * world)(*[table.data_ptr() + i * share * W for i in range(world)])
        nbs = (ctypes.c_uint64 * world)(*[share * W] * world)
        timed("a2a_copy (contiguous, interleaved segments; remote bytes counted)",
              lambda: _cabi.check(lib.xchg_copy_segments(world, srcs, outs, nbs, stream)), remote_bytes)
        libc = ctypes.CDLL(None)
        libc.setenv(b"XCHG_COPY_BULK", b"1", 1)
        timed("a2a_copy_tma (cp.async.bulk, interleaved segments; remote bytes counted)",
              lambda: _cabi.check(lib.xchg_copy_segments(world, srcs, outs, nbs, stream)), remote_bytes)
        libc.setenv(b"XCHG_COPY_BULK", b"0", 1)
        # copy engines: one cudaMemcpyAsync per peer, each on its own stream
        ce_streams = [torch.cuda.Stream(device=rt.device) for _ in range(world)]
        views = [hdl.get_buffer(d, (share * W,), torch.uint8, land_off + rank * share * W) for d in order]

        def a2a_ce():
            cur = torch.cuda.current_stream(rt.device)
            for i, d in enumerate(order):
                if d == rank:
                    continue
                ce_streams[i].wait_stream(cur)
                with torch.cuda.stream(ce_streams[i]):
                    views[i].copy_(table[i * share * W:(i + 1) * share * W], non_blocking=True)
            for i, d in enumerate(order):
                if d != rank:
                    cur.wait_stream(ce_streams[i])

        timed("a2a_ce (cudaMemcpyAsync per peer on its own stream; remote bytes counted)", a2a_ce, remote_bytes)
        send = table[: share * W * world]
        recv = local_out[: share * W * world]
        timed("a2a_nccl all_to_all_single (remote bytes counted)",
              lambda: dist.all_to_all_single(recv, send), remote_bytes)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

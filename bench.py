#!/usr/bin/env python
"""Benchmark of the exchange hot path: Dataset.sort on GraySort rows.

  python bench.py [--gpus N] [--steps K] [--warmup W]          # this engine
  python bench.py --impl reference [--steps K] [--warmup W]   # CPU reference arm

Workload (BASELINE.json configs[1]): int64 key + 90-byte binary payload rows,
125 M rows per GPU (12.25 GB) in 12 blocks -- 1 B rows at 8 GPUs, weak scaling.
The 1 B-row config does not fit one GPU (input + output = 196 GB > 180 GB), so
N=1 runs its per-GPU shard; ``config.workload`` says so.

One step = one full ``Dataset.sort('c_0')`` of the resident blocks: boundary
sampling (K1), key normalisation + histograms (K2), onesweep radix passes (K3),
boundary search (K4), key + payload gather (K5) [+ exchange and merge at N>1].
``value`` has the inputs resident in HBM; ``e2e`` goes through the public
host-buffer API (``DeviceExchange.sort_blocks``) from pinned host Arrow blocks
and back, copies inside the timed region.  Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import statistics
import subprocess
import sys
import tempfile
import threading
import time

# (no expandable_segments: with it the allocator re-mapped the ~12 GB output buffer on some
# ranks every step, ~1 ms per GB of host stall, and the 8-GPU step waited for those ranks;
# Runtime.empty rounds large buffers to sticky size classes instead)

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Dataset.sort rows/s (GraySort rows: int64 key + 90 B payload)"
UNIT = "rows/s"
ROW_BYTES = 98
PAYLOAD = 90

KERNEL_NAMES = {0: "k_normalize_hist", 1: "k_scan_hist", 2: "k_onesweep", 3: "k_gather_rows",
                4: "k_gather_fixed", 5: "k_gather_bitmap", 6: "k_lengths_scan", 7: "k_partition_bounds",
                8: "k_offsets_*", 9: "misc", 10: "k_merge", 11: "k_shuffle", 12: "k_aggregate", 13: "k_local_sort", 14: "k_msd_aux"}
# algorithmic bytes per row per launch
KERNEL_BYTES_PER_ROW = {0: 16, 2: 24, 3: 2 * PAYLOAD, 13: 24}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def exchange_roofline(total_rows: int, world: int, hbm_gbs: float, ms_per_step: float) -> dict:
    """lower bound of one step = the slower of the HBM bound (every row read
    once and written once: 2 x row bytes) and the NVLink bound (a row leaves its GPU with
    probability (G-1)/G; 900 GB/s per direction per GPU nominal, 770 GB/s measured peer copy,
    B200_PROFILING.md); fraction = bound / measured step."""
    per_gpu_bytes = total_rows * ROW_BYTES / world
    t_hbm = 2 * per_gpu_bytes / (hbm_gbs * 1e9) * 1e3
    out = {"hbm_bound_ms": t_hbm}
    for name, gbs in (("nominal_900", 900.0), ("measured_770", 770.0)):
        t_nv = per_gpu_bytes * (world - 1) / world / (gbs * 1e9) * 1e3
        lb = max(t_hbm, t_nv)
        out[name] = {"nvlink_bound_ms": t_nv, "bound_ms": lb,
                     "binding": "nvlink" if t_nv > t_hbm else "hbm", "frac": lb / ms_per_step}
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        if shutil.which("nvidia-smi") is None:
            return
        fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
        os.close(fd)
        self.proc = subprocess.Popen(
            ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
             "-i", str(self.idx)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
                for nm, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        finally:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm))
        return out


# ----------------------------------------------------------------------------
# CPU reference arm (oracle port of the reference's pyarrow/numpy exchange)
# ----------------------------------------------------------------------------

def _shm_dir():
    base = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > (32 << 30) else None
    return tempfile.mkdtemp(prefix="xchg_cpu_", dir=base)


def _write_ipc(path, table):
    import pyarrow as pa

    with pa.OSFile(path, "wb") as f:
        with pa.ipc.new_file(f, table.schema) as w:
            w.write_table(table)


def _read_ipc(path):
    import pyarrow as pa

    return pa.ipc.open_file(pa.memory_map(path, "r")).read_all()


def _cpu_map_task(args):
    """One reference map task (SortTaskSpec.map): block in shared memory -> R sorted
    partitions in shared memory (as plasma objects would be)."""
    i, in_path, out_dir, boundaries, null_scan = args
    from oracle import exchange_oracle as O

    blk = _read_ipc(in_path)
    parts = O.sort_and_partition(blk, boundaries, ["c_0"], [False], null_scan=null_scan)
    paths = []
    for j, p in enumerate(parts):
        path = os.path.join(out_dir, f"m{i}_{j}.arrow")
        _write_ipc(path, p)
        paths.append(path)
    return paths


def _cpu_reduce_task(args):
    j, paths, out_dir = args
    from oracle import exchange_oracle as O

    out = O.merge_sorted_blocks([_read_ipc(p) for p in paths], ["c_0"], [False])
    path = os.path.join(out_dir, f"r{j}.arrow")
    _write_ipc(path, out)
    return out.num_rows


def make_cpu_blocks(num_blocks: int, rows_per_block: int, seed: int = 2):
    import numpy as np
    import pyarrow as pa

    rng = np.random.default_rng(seed)
    blocks = []
    offs = np.arange(rows_per_block + 1, dtype=np.int32) * PAYLOAD
    for _ in range(num_blocks):
        keys = rng.integers(0, 2**63 - 1, rows_per_block, dtype=np.int64)
        payload = rng.integers(0, 256, rows_per_block * PAYLOAD, dtype=np.uint8)
        c1 = pa.Array.from_buffers(pa.binary(), rows_per_block,
                                   [None, pa.py_buffer(offs), pa.py_buffer(payload)])
        blocks.append(pa.table({"c_0": keys, "c_1": c1}))
    return blocks


class CpuReference:
    """The reference's CPU exchange on a bounded sample: the pull-based schedule
    (all SortTaskSpec.map tasks, then all .reduce tasks --
    pull_based_shuffle_task_scheduler.py:84-120) on a pool of one process per core,
    blocks and partitions handed over as Arrow IPC in shared memory (what Ray's
    plasma store does).  ``null_scan=True`` is the reference as-is (incl. the
    O(R*N) ``col == None`` scan of find_partition_index, util.py:769)."""

    def __init__(self, cores: int, rows_per_block: int, num_blocks: int):
        import multiprocessing as mp

        self.cores = cores
        self.rows_per_block = rows_per_block
        self.num_blocks = num_blocks
        self.dir = _shm_dir()
        blocks = make_cpu_blocks(num_blocks, rows_per_block)
        self.in_paths = []
        for i, b in enumerate(blocks):
            p = os.path.join(self.dir, f"in{i}.arrow")
            _write_ipc(p, b)
            self.in_paths.append(p)
        from oracle import exchange_oracle as O
        import numpy as np

        # driver-side boundary sampling (sort_task_spec.py:154-226), untimed setup
        rng = np.random.default_rng(0)
        n_s = O.samples_per_block(num_blocks, num_blocks)
        samples = []
        for b in blocks:
            idx = rng.choice(b.num_rows, size=min(n_s, b.num_rows), replace=False)
            samples += [(int(k),) for k in b["c_0"].to_numpy()[idx]]
        self.boundaries = O.boundaries_from_samples(samples, [False], num_blocks)
        del blocks
        self.pool = mp.get_context("fork").Pool(cores)
        self.total_rows = rows_per_block * num_blocks

    def step(self, null_scan: bool = True) -> float:
        out_dir = tempfile.mkdtemp(prefix="step_", dir=self.dir)
        t0 = time.perf_counter()
        maps = self.pool.map(_cpu_map_task,
                             [(i, p, out_dir, self.boundaries, null_scan) for i, p in enumerate(self.in_paths)],
                             chunksize=1)
        rows = self.pool.map(_cpu_reduce_task,
                             [(j, [m[j] for m in maps], out_dir) for j in range(self.num_blocks)],
                             chunksize=1)
        dt = time.perf_counter() - t0
        assert sum(rows) == self.total_rows
        shutil.rmtree(out_dir, ignore_errors=True)
        return dt

    def close(self):
        self.pool.close()
        self.pool.join()
        shutil.rmtree(self.dir, ignore_errors=True)


def run_reference(args):
    """--impl reference: rank 0 only; other ranks exit 0 without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    rows_per_block = args.cpu_rows_per_block
    nblocks = args.cpu_blocks or cores
    ref = CpuReference(cores, rows_per_block, nblocks)
    try:
        for _ in range(args.warmup):
            ref.step()
        times = [ref.step() for _ in range(args.steps)]
        prim = ref.step(null_scan=False)
    finally:
        ref.close()
    t = sum(times) / len(times)
    value = ref.total_rows / t
    sample = (f"{nblocks} blocks x {rows_per_block} GraySort rows ({ref.total_rows * ROW_BYTES / 1e9:.2f} GB), "
              f"R={nblocks}, pull-based map/reduce on {cores} processes, Arrow IPC hand-off in shared memory")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": config_dict(args, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "scale": ref.total_rows / float(args.rows_per_gpu * max(1, args.gpus)),
                         "gb_per_s": value * 2 * ROW_BYTES / 1e9,
                         "primitives_only_value": ref.total_rows / prim,
                         "note": "oracle port = the reference's own pyarrow/numpy call sequence; "
                                 "value is the reference as-is, primitives_only_value drops its "
                                 "O(R*N) `col == None` scan"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------

def config_dict(args, world):
    """The same dict in both arms (the driver compares them)."""
    return {"workload": workload_name(args), "rows_total": args.rows_per_gpu * world,
            "cache": "inputs (12.25 GB/GPU) far exceed the 126 MB L2; no flush needed",
            "partitioning": "range (sampled boundaries), R = number of input blocks"}


def workload_name(args):
    return (f"GraySort-style Dataset.sort: int64 key + 90 B binary payload, {args.rows_per_gpu} rows/GPU "
            f"in {args.blocks_per_gpu} blocks ({args.rows_per_gpu * ROW_BYTES / 1e9:.2f} GB/GPU; "
            f"BASELINE.json configs[1] = 1e9 rows at 8 GPUs, weak-scaled shard at fewer GPUs because "
            f"1e9 rows x 98 B in + out exceeds one GPU's 180 GB)")


def build_resident_dataset(rt, n, blocks, seed):
    """Synthetic GraySort rows generated directly in HBM (seeded)."""
    import pyarrow as pa
    import torch

    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceDataset

    g = torch.Generator(device=rt.device)
    g.manual_seed(seed)
    keys = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=rt.device, generator=g)
    data = rt.empty(n * PAYLOAD)
    chunk = 1 << 28
    for s in range(0, n * PAYLOAD, chunk):
        e = min(n * PAYLOAD, s + chunk)
        data[s:e] = torch.randint(0, 256, (e - s,), dtype=torch.uint8, device=rt.device, generator=g)
    # bytes 0..7 of every payload row repeat the row's key: after the sort a row that lost
    # its payload (or a payload that lost its key) is detectable (``verify_output``)
    data[: n * PAYLOAD].view(n, PAYLOAD)[:, :8].copy_(keys.view(torch.uint8).view(n, 8))
    schema = pa.schema([("c_0", pa.int64()), ("c_1", pa.binary())])
    c0 = dev.DeviceColumn(pa.int64(), n, "fixed", 8, keys.view(torch.uint8))
    c1 = dev.DeviceColumn(pa.binary(), n, "var", 1, data, None, None, 0, uniform=PAYLOAD)
    table = dev.DeviceTable(["c_0", "c_1"], [c0, c1], n, schema)
    per = n // blocks
    ranges = [(i * per, (i + 1) * per if i < blocks - 1 else n) for i in range(blocks)]
    return DeviceDataset(table, ranges), keys, data


def host_blocks_from_device(rt, keys, data, ranges):
    """Pinned-host Arrow blocks holding the same rows (e2e input)."""
    import numpy as np
    import pyarrow as pa
    import torch

    n = keys.numel()
    keys_h = torch.empty(n, dtype=torch.int64, pin_memory=True)
    keys_h.copy_(keys)
    data_h = torch.empty(n * PAYLOAD, dtype=torch.uint8, pin_memory=True)
    data_h.copy_(data[: n * PAYLOAD])
    max_rows = max(hi - lo for lo, hi in ranges)
    offs_h = torch.empty(max_rows + 1, dtype=torch.int32, pin_memory=True)
    offs_h.copy_(torch.arange(max_rows + 1, dtype=torch.int32) * PAYLOAD)
    torch.cuda.synchronize()
    kbuf, dbuf, obuf = pa.py_buffer(keys_h.numpy()), pa.py_buffer(data_h.numpy()), pa.py_buffer(offs_h.numpy())
    blocks = []
    for lo, hi in ranges:
        m = hi - lo
        c0 = pa.Array.from_buffers(pa.int64(), m, [None, kbuf.slice(lo * 8, m * 8)])
        c1 = pa.Array.from_buffers(pa.binary(), m, [None, obuf.slice(0, (m + 1) * 4),
                                                    dbuf.slice(lo * PAYLOAD, m * PAYLOAD)])
        blocks.append(pa.table({"c_0": c0, "c_1": c1}))
    return blocks, (keys_h, data_h, offs_h)


def _wrapping_sum(t):
    """Sum of an int64 tensor modulo 2^64 (torch's integer sum wraps)."""
    return t.sum(dtype=t.dtype)


def input_fingerprint(keys, data, n):
    """Order-independent fingerprint of this rank's rows: count, sum of keys and sum of the
    payload word at bytes 8..15 of every row, both modulo 2^64."""
    import torch

    w = data[: n * PAYLOAD].view(n, PAYLOAD)[:, 8:16].contiguous().view(torch.int64).view(n)
    return torch.stack([torch.tensor(n, dtype=torch.int64, device=keys.device), _wrapping_sum(keys),
                        _wrapping_sum(w)])


def verify_output(out, fp_in, world):
    """Checks on the LAST timed step's output, every rank, on the device:
    order            keys non-decreasing inside the rank and block ranges tile its rows;
    rank_boundaries  last key of rank g <= first key of the next rank that holds rows;
    checksum         row count, sum of keys and sum of a payload word (mod 2^64), summed
                     over ranks, equal the input's;
    payload_keys     bytes 0..7 of every payload row equal the key beside it."""
    import torch
    import torch.distributed as dist

    dev_ = fp_in.device
    m = out.num_rows
    if m:
        k = out.table.column("c_0").data[: m * 8].view(torch.int64)
        p = out.table.column("c_1").data[: m * PAYLOAD].view(m, PAYLOAD)
        order = bool((k[1:] >= k[:-1]).all().item()) if m > 1 else True
        pk = bool((p[:, :8].contiguous().view(torch.int64).view(m) == k).all().item())
        w = p[:, 8:16].contiguous().view(torch.int64).view(m)
        fp_out = torch.stack([torch.tensor(m, dtype=torch.int64, device=dev_), _wrapping_sum(k), _wrapping_sum(w)])
        edge = torch.stack([k[0], k[-1], torch.tensor(1, dtype=torch.int64, device=dev_)])
    else:
        order, pk = True, True
        fp_out = torch.zeros(3, dtype=torch.int64, device=dev_)
        edge = torch.zeros(3, dtype=torch.int64, device=dev_)
    last = 0
    for lo, hi in out.block_ranges:
        if hi > lo:
            order = order and lo == last
            last = hi
    order = order and last == m
    fps = torch.stack([fp_in, fp_out])
    flags = torch.tensor([int(order), int(pk)], dtype=torch.int64, device=dev_)
    if world > 1:
        dist.all_reduce(fps)  # int64 sums wrap: still a valid fingerprint mod 2^64
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        edges = [torch.empty_like(edge) for _ in range(world)]
        dist.all_gather(edges, edge)
    else:
        edges = [edge]
    es = [e.tolist() for e in edges]
    es = [e for e in es if e[2]]
    rb = all(es[i][1] <= es[i + 1][0] for i in range(len(es) - 1))
    order, pk = bool(flags[0].item()), bool(flags[1].item())
    return {"order": order, "rank_boundaries": rb, "checksum": bool((fps[0] == fps[1]).all().item()),
            "payload_keys": pk, "rows": int(fps[1][0].item())}


def run_b200(args):
    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and local_rank != 0:
        time.sleep(0.5)  # let local rank 0 win the (idempotent) build check
    ge.build()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from cluster_anywhere_b200 import SortKey, _cabi
    from cluster_anywhere_b200 import device as dev
    from cluster_anywhere_b200.exchange import DeviceExchange

    lib = _cabi.load()
    rt = dev.Runtime()
    ex = DeviceExchange(rt, trace=args.trace)
    n = args.rows_per_gpu
    ds, keys, data = build_resident_dataset(rt, n, args.blocks_per_gpu, seed=2 + rank)
    sk = SortKey("c_0")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(i):
        if world > 1:
            out = ex.sort_distributed(ds, sk, seed=i)
        else:
            out = ex.sort(ds, sk, seed=i)
        return out

    # ---- device-resident timing ----
    for i in range(args.warmup):
        out = one_step(i)
        del out
    barrier()
    ex.phase_ms = {}
    ex.timeline_ms = {}
    ex.gpu_marks = []
    import resource as _resource
    ru0 = _resource.getrusage(_resource.RUSAGE_SELF)
    if args.trace:
        from cluster_anywhere_b200.exchange import collective as coll

        coll.TRACE = ex.phase_ms
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    lib.xchg_profile_enable(1)
    launches0 = lib.xchg_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    alloc_log = []
    for i in range(args.steps):
        out = one_step(100 + i)
        if i < args.steps - 1:
            del out
        if getattr(ex, "timeline", False):
            st = torch.cuda.memory_stats()
            alloc_log.append((st.get("num_device_alloc", 0), st.get("num_device_free", 0),
                              round(torch.cuda.memory_reserved() / 1e9, 2)))
    ev1.record()
    barrier()
    launches = lib.xchg_launch_count() - launches0
    lib.xchg_profile_enable(0)
    clocks = sampler.stop() if rank == 0 else None
    ms = ev0.elapsed_time(ev1)
    tmax = torch.tensor([ms], dtype=torch.float64, device=rt.device)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_per_step = float(tmax.item()) / args.steps
    total_rows = n * world
    verified = verify_output(out, input_fingerprint(keys, data, n), world)
    del out
    if args.no_verify:
        verified["skipped"] = "--no-verify (a timing experiment; this line is not a bench value)"
    elif not (verified["order"] and verified["rank_boundaries"] and verified["checksum"]
            and verified["payload_keys"] and verified["rows"] == total_rows):
        if rank == 0:
            print(json.dumps({"metric": METRIC, "error": "output verification failed", "verified": verified,
                              "n_gpus": world}), flush=True)
        raise SystemExit(3)
    value = total_rows / (ms_per_step / 1e3)

    # per-kernel device time over the timed region (rank 0's GPU)
    import ctypes

    kernels = []
    for kid, name in KERNEL_NAMES.items():
        tot, cnt = ctypes.c_double(), ctypes.c_uint64()
        _cabi.check(lib.xchg_profile_read(kid, ctypes.byref(tot), ctypes.byref(cnt)))
        if cnt.value:
            kernels.append({"id": kid, "kernel": name, "launches_per_step": cnt.value / args.steps,
                            "ms_per_step": tot.value / args.steps, "avg_launch_ms": tot.value / cnt.value})
    kernels.sort(key=lambda k: -k["ms_per_step"])
    peak, peak_src = peaks()
    roofline = None
    for k in kernels:
        if k["id"] in KERNEL_BYTES_PER_ROW:
            # rows one step sends through this kernel class on this rank: every normalise /
            # onesweep launch covers the n resident rows; the local sorts of a step cover
            # them once in total (one launch per segment-size class); the row gather covers
            # them once, or twice when distributed (into the peers' arenas, one launch per
            # destination, then the final take)
            if k["id"] == 13:
                rows_per_step = n
            elif k["id"] == 3:
                rows_per_step = n * (2 if world > 1 else 1)
            else:
                rows_per_step = n * k["launches_per_step"]
            bpl = KERNEL_BYTES_PER_ROW[k["id"]] * rows_per_step / k["launches_per_step"]
            k["algorithmic_bytes_per_launch"] = bpl
            k["achieved_gbs"] = bpl / (k["avg_launch_ms"] * 1e-3) / 1e9
            k["frac_of_peak"] = k["achieved_gbs"] / peak
    dom = next((k for k in kernels if "achieved_gbs" in k), None)
    if dom is not None:
        roofline = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved_gbs"], "peak": peak,
                    "unit": "GB/s", "frac": dom["achieved_gbs"] / peak, "traffic": None,
                    "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_launch"],
                    "share_of_step": dom["ms_per_step"] / ms_per_step}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr) and world == 1:  # captured per launch at one GPU (n rows per launch)
            try:
                with open(tr) as f:
                    roofline["traffic"] = json.load(f).get(dom["kernel"])
            except Exception:
                pass

    # ---- end to end through the host-buffer API (pinned Arrow blocks in and out) ----
    e2e = None
    if not args.no_e2e:
        blocks, keep = host_blocks_from_device(rt, keys, data, ds.block_ranges)
        h2d = sum(b.nbytes for b in blocks)

        breakdown = {} if os.environ.get("XCHG_E2E_BREAKDOWN", "0") == "1" else None

        def e2e_step(i):
            if breakdown is not None:
                # diagnostics: the three phases one after the other, a synchronise between them
                t = [time.perf_counter()]
                dsx = ex.stage(blocks)
                torch.cuda.synchronize()
                t.append(time.perf_counter())
                outx = ex.sort_distributed(dsx, sk, seed=i) if world > 1 else ex.sort(dsx, sk, seed=i)
                torch.cuda.synchronize()
                t.append(time.perf_counter())
                res = ex.export(outx)
                t.append(time.perf_counter())
                for nm, a, b in (("stage_h2d", 0, 1), ("sort", 1, 2), ("export_d2h", 2, 3)):
                    breakdown.setdefault(nm, []).append(round((t[b] - t[a]) * 1e3, 1))
                return res
            if world > 1:
                return ex.sort_blocks_distributed(blocks, sk, seed=i)
            return ex.sort_blocks(blocks, sk, seed=i)

        for i in range(max(1, min(args.warmup, 2))):
            outs = e2e_step(i)
            del outs
        e_steps = max(1, min(args.steps, args.e2e_steps))
        barrier()
        t0 = time.perf_counter()
        d2h = 0
        for i in range(e_steps):
            outs = e2e_step(100 + i)
            d2h = sum(o.nbytes for o in outs)
            rows = sum(o.num_rows for o in outs)
            del outs
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e_steps
        agg = torch.tensor([dt, -float(rows), -float(h2d), -float(d2h)], dtype=torch.float64, device=rt.device)
        if world > 1:
            # slowest rank's time; rows and bytes summed over ranks
            tm = agg[:1].clone()
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            sums = -agg[1:].clone()
            dist.all_reduce(sums)
            dt, rows, h2d, d2h = float(tm.item()), int(sums[0].item()), int(sums[1].item()), int(sums[2].item())
        assert rows == total_rows
        if breakdown is not None:
            print(f"[e2e breakdown rank {rank}] ms per call (warm-up calls first): {json.dumps(breakdown)}",
                  file=sys.stderr, flush=True)
        e2e = {"value": total_rows / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": dt * 1e3, "steps": e_steps,
               "api": ("DeviceExchange.sort_blocks_distributed" if world > 1 else "DeviceExchange.sort_blocks")
                      + "(pinned host Arrow blocks) -> host Arrow blocks; wall clock, max over ranks"}
        del blocks, keep

    # ---- the path's other operators (N=1 only): random_shuffle, groupby.aggregate, string-key
    # sort and configs[0], same event / per-kernel bookkeeping (scripts/bench_ops.py) ----
    ops = None
    if world == 1 and not args.no_ops:
        del ds, keys, data
        ds = keys = data = None
        torch.cuda.empty_cache()
        import importlib.util

        spec = importlib.util.spec_from_file_location("bench_ops", os.path.join(ROOT, "scripts", "bench_ops.py"))
        bench_ops = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(bench_ops)
        ops = bench_ops.run_ops(rt, ex, lib, peak, steps=5, warmup=3)

    # ---- CPU baseline (rank 0, N=1 only): bounded sample of the same workload ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        ds = keys = data = None
        torch.cuda.empty_cache()
        cores = len(os.sched_getaffinity(0))
        ref = CpuReference(cores, args.cpu_rows_per_block, args.cpu_blocks or cores)
        try:
            t_cpu = ref.step()
            t_prim = ref.step(null_scan=False)
        finally:
            ref.close()
        sample = (f"{ref.num_blocks} blocks x {ref.rows_per_block} GraySort rows "
                  f"({ref.total_rows * ROW_BYTES / 1e9:.2f} GB), R={ref.num_blocks}, pull-based map/reduce on "
                  f"{cores} processes, Arrow IPC hand-off in shared memory")
        cpu = {"value": ref.total_rows / t_cpu, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": sample, "scale": ref.total_rows / float(total_rows),
               "gb_per_s": ref.total_rows / t_cpu * 2 * ROW_BYTES / 1e9,
               "primitives_only_value": ref.total_rows / t_prim}

    if getattr(ex, "timeline", False):
        torch.cuda.synchronize()
        ru1 = _resource.getrusage(_resource.RUSAGE_SELF)
        print(f"[host rank {rank}] cpus={len(os.sched_getaffinity(0))} loadavg={os.getloadavg()} "
              f"utime={ru1.ru_utime - ru0.ru_utime:.3f}s stime={ru1.ru_stime - ru0.ru_stime:.3f}s "
              f"vol_cs={ru1.ru_nvcsw - ru0.ru_nvcsw} invol_cs={ru1.ru_nivcsw - ru0.ru_nivcsw}",
              file=sys.stderr, flush=True)
        print(f"[alloc rank {rank}] (device allocs, device frees, reserved GB) after each timed step: {alloc_log}",
              file=sys.stderr, flush=True)
        print(f"[gpu timeline rank {rank}] mean ms from step begin (timed steps + e2e steps): "
              + json.dumps(ex.gpu_timeline()), file=sys.stderr, flush=True)
        print(f"[timeline rank {rank}] host ms between stamps, summed over the timed steps (+ e2e steps if run): "
              + json.dumps({k: round(v, 2) for k, v in ex.timeline_ms.items()}), file=sys.stderr, flush=True)
    if args.trace:
        print(f"[trace rank {rank}] phase ms over the {args.steps} timed steps (e2e.*: over warm-up + e2e steps): "
              + json.dumps({k: round(v, 2) for k, v in ex.phase_ms.items()}), file=sys.stderr, flush=True)
    xr = exchange_roofline(total_rows, world, peak, ms_per_step)
    if world > 1 and roofline is not None:
        # the step's binding bound at N > 1 is the NVLink egress of every GPU:
        # bytes that must leave a GPU / step time, against 900 GB/s per direction
        link_bytes = n * ROW_BYTES * (world - 1) / world
        roofline = {"bound": "nvlink", "achieved": link_bytes / (ms_per_step * 1e-3) / 1e9, "peak": 900.0,
                    "unit": "GB/s", "frac": xr["nominal_900"]["frac"], "traffic": None,
                    "peak_source": "NVLink 5 nominal, per direction per GPU (B200_PROFILING.md)",
                    "algorithmic_bytes_per_step_per_gpu": link_bytes,
                    "dominant_kernel_hbm": roofline}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": config_dict(args, world),
            "gb_per_s_algorithmic": value * 2 * ROW_BYTES / 1e9,
            "frac_of_hbm_roofline": value * 2 * ROW_BYTES / 1e9 / (peak * world),
            "exchange_roofline": xr,
            "gpu_launches": int(launches),
            "verified": verified,
            "clocks": clocks,
            "roofline": roofline,
            "kernels": kernels,
            "e2e": e2e,
            "cpu_baseline": cpu,
            "ops": ops,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows-per-gpu", type=int, default=125_000_000)
    ap.add_argument("--blocks-per-gpu", type=int, default=12)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-verify", action="store_true",
                    help="timing experiments with XCHG_DEBUG_SKIP: do not abort on a wrong output")
    ap.add_argument("--no-ops", action="store_true", help="skip the secondary operators' lines (ops key)")
    ap.add_argument("--trace", action="store_true",
                    help="diagnostics: synchronise after every phase of the distributed sort and print "
                         "per-phase wall-clock ms (serialises the pipeline; not a bench number)")
    ap.add_argument("--cpu-rows-per-block", type=int, default=2_000_000)
    ap.add_argument("--cpu-blocks", type=int, default=0, help="default: one block per host core")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()

"""What the multi-GPU exchange needs from "the other ranks", behind one small interface.

``TorchComm`` is production: one process per GPU, ``torch.distributed`` (NCCL) for the
few small collectives and ``torch.distributed._symmetric_memory`` for the peer-mapped
arenas -- plumbing; every byte of data is moved by this engine's own kernels.

``ThreadComm`` runs G ranks as G threads of ONE process on ONE device: each rank has its
own streams, its own "arena" (a plain device buffer whose address the others know) and
the same kernels store into it.  It exists so that the N > 1 code -- arena layouts with
non-zero source prefixes, chunked stores, arrival flags -- is exercised by the ``-m gpu``
tests on a single-GPU box (and, over the host-memory test double, by the CPU tests).

In the reference the ranks are Ray tasks and the "communicator" is the object store
(python/ray/data/_internal/planner/exchange/pull_based_shuffle_task_scheduler.py:84-120).
"""
from __future__ import annotations

import threading
from typing import List, Optional

import torch

from .. import device as dev


class PeerArena:
    """One byte buffer per rank, addressable by every rank."""

    def __init__(self, nbytes: int, buf: torch.Tensor, ptrs: List[int], keepalive=None):
        self.nbytes = int(nbytes)
        self.buf = buf
        self.ptrs = [int(p) for p in ptrs]
        self.hdl = keepalive


class TorchComm:
    def __init__(self, group=None):
        import torch.distributed as dist

        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.virtual = False

    def all_gather(self, t: torch.Tensor) -> torch.Tensor:
        """[world, len(t)] on the device; stream-ordered on the current stream."""
        import torch.distributed as dist

        out = torch.empty(self.world * t.numel(), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t.contiguous(), group=self.group)
        return out.view(self.world, t.numel())

    def all_gather_object(self, obj) -> list:
        import torch.distributed as dist

        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def barrier(self):
        import torch.distributed as dist

        dist.barrier(group=self.group)

    def fence(self, rt: dev.Runtime):
        """Stream-ordered: when it completes every rank's earlier stores are visible."""
        import torch.distributed as dist

        flag = torch.zeros(1, dtype=torch.int32, device=rt.device)
        dist.all_reduce(flag, group=self.group)

    def new_arena(self, rt: dev.Runtime, nbytes: int) -> PeerArena:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        buf = symm_mem.empty(int(nbytes), dtype=torch.uint8, device=rt.device)
        hdl = symm_mem.rendezvous(buf, group=self.group if self.group is not None else dist.group.WORLD)
        return PeerArena(nbytes, buf, [int(p) for p in hdl.buffer_ptrs], keepalive=hdl)


class _ThreadWorld:
    def __init__(self, world: int):
        self.world = world
        self.bar = threading.Barrier(world)
        self.slots: List[object] = [None] * world
        self.failed = False


class ThreadComm:
    """Rank ``rank`` of a ``ThreadWorld`` (see ``run_virtual_world``)."""

    def __init__(self, shared: _ThreadWorld, rank: int):
        self._w = shared
        self.rank = rank
        self.world = shared.world
        self.group = None
        self.virtual = True

    def _wait(self):
        try:
            self._w.bar.wait(timeout=120)
        except threading.BrokenBarrierError:
            raise RuntimeError("a rank of the virtual world failed") from None

    def _exchange(self, obj) -> list:
        self._w.slots[self.rank] = obj
        self._wait()
        out = list(self._w.slots)
        self._wait()
        return out

    def all_gather(self, t: torch.Tensor) -> torch.Tensor:
        if t.is_cuda:
            torch.cuda.current_stream(t.device).synchronize()
        parts = self._exchange(t)
        out = torch.stack([p.reshape(-1) for p in parts])
        if t.is_cuda:
            # the other ranks may reuse their buffers as soon as the second barrier of the
            # NEXT exchange is passed: finish the copy first
            torch.cuda.current_stream(t.device).synchronize()
        self._wait()
        return out

    def all_gather_object(self, obj) -> list:
        return self._exchange(obj)

    def barrier(self):
        self._wait()

    def fence(self, rt: dev.Runtime):
        rt.synchronize()
        self._wait()

    def new_arena(self, rt: dev.Runtime, nbytes: int) -> PeerArena:
        buf = torch.empty(int(nbytes), dtype=torch.uint8, device=rt.device)
        ptrs = self._exchange(buf.data_ptr())
        return PeerArena(nbytes, buf, ptrs)


def as_comm(group=None):
    """A process group (or None = the default group) -> TorchComm; a Comm stays itself;
    no process group at all -> a one-rank world."""
    if isinstance(group, (TorchComm, ThreadComm)):
        return group
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return TorchComm(group)
    return ThreadComm(_ThreadWorld(1), 0)  # no process group: a world of one


def run_virtual_world(world: int, fn, device: Optional[torch.device] = None) -> list:
    """Run ``fn(comm, rank)`` on ``world`` threads, one virtual rank each, all on one device
    (every thread gets its own CUDA stream).  Returns the ranks' results; re-raises the
    first failure."""
    shared = _ThreadWorld(world)
    results: List[object] = [None] * world
    errors: List[Optional[BaseException]] = [None] * world
    use_cuda = device is not None and torch.device(device).type == "cuda"

    def body(r: int):
        try:
            comm = ThreadComm(shared, r)
            if use_cuda:
                torch.cuda.set_device(device)
                with torch.cuda.stream(torch.cuda.Stream(device=device)):
                    results[r] = fn(comm, r)
                    torch.cuda.current_stream(device).synchronize()
            else:
                results[r] = fn(comm, r)
        except BaseException as e:  # noqa: BLE001 - re-raised below
            errors[r] = e
            shared.bar.abort()

    threads = [threading.Thread(target=body, args=(r,), name=f"vrank{r}") for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    first = next((e for e in errors if e is not None and not isinstance(e, RuntimeError)), None)
    first = first or next((e for e in errors if e is not None), None)
    if first is not None:
        raise first
    return results

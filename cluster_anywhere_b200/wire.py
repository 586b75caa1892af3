"""Wire format <-> device staging: the step either side of the exchange.

Between tasks the reference moves blocks through plasma as pickled Arrow tables: per-chunk
buffer payloads for most types and, as the general fallback, one Arrow IPC STREAM per table
(``_arrow_table_ipc_reduce`` / ``_restore_table_from_ipc``,
python/ray/_private/arrow_serialization.py:785-809; registered for ``pa.Table`` at :103-118).
A reader of such a stream gets record batches that are zero-copy views of the sender's bytes.

This module lets the engine take those bytes as they are:

* :func:`table_from_ipc` opens a stream (or file) held in ANY host buffer -- ``bytes``, a
  ``memoryview`` of a shared-memory segment, a ``pa.Buffer`` -- without copying;
* :class:`HostRegistration` page-locks caller-owned host memory (``cudaHostRegister``) so that
  the host->device copies out of it are real asynchronous DMA instead of the driver's
  synchronous staged copy of pageable memory.  Page-locking costs about as much as one copy
  of the range, so it pays for memory that is reused -- an object-store arena registered
  once -- not per call; :func:`stage_ipc` therefore registers only on request;
* :func:`stage_ipc` = open + ``DeviceExchange.stage`` (blocks stay views of the wire bytes
  until the copy engine reads them: int32 offsets are widened, slices re-based and chunks
  concatenated on the device, ``device.stage_column``);
* :func:`export_ipc` writes every output block as one IPC stream into pinned host memory
  (multi-threaded memcpy of the already pinned column buffers).

Nothing here computes on the CPU: it is framing and address bookkeeping.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import pyarrow as pa
import torch

from . import _cabi
from . import device as dev



def table_to_ipc(table: pa.Table) -> pa.Buffer:
    """One Arrow IPC stream holding ``table`` (the writer of arrow_serialization.py:792-801)."""
    sink = pa.BufferOutputStream()
    with pa.ipc.new_stream(sink, table.schema) as w:
        w.write_table(table)
    return sink.getvalue()


def table_from_ipc(buf) -> pa.Table:
    """Zero-copy view of the table inside an IPC stream or file held in ``buf``
    (arrow_serialization.py:804-809 reads the stream the same way)."""
    b = buf if isinstance(buf, pa.Buffer) else pa.py_buffer(buf)
    if b.size >= 6 and b.slice(0, 6).to_pybytes() == b"ARROW1":
        return pa.ipc.open_file(b).read_all()
    return pa.ipc.open_stream(b).read_all()


class HostRegistration:
    """Page-locked windows over caller-owned host memory.  ``add`` is idempotent for ranges
    already registered (here or by someone else); ``release`` undoes what this object did.
    On a runtime without CUDA (the tests' host-memory double) it does nothing."""

    def __init__(self, rt: Optional[dev.Runtime] = None):
        self._rt = rt or dev.default_runtime()
        self._ranges: List[int] = []
        self.registered_bytes = 0
        self.failed = 0

    def add(self, *bufs) -> int:
        """Page-lock the buffers (anything ``pa.py_buffer`` accepts); overlapping ranges are
        registered as one.  The ranges are registered byte-exact, NOT rounded out to pages: a
        heap neighbour that shares a buffer's first or last page must stay "pageable" in the
        driver's eyes, or a later copy into it -- half inside a registered range -- is refused
        (seen as `invalid argument` on a plain ``tensor.cpu()``).  -> number of ranges from
        which copies are now asynchronous."""
        if not self._rt.is_cuda:
            return 0
        spans = []
        for buf in bufs:
            b = buf if isinstance(buf, pa.Buffer) else pa.py_buffer(buf)
            if b.size:
                spans.append((b.address, b.address + b.size))
        spans.sort()
        merged: List[List[int]] = []
        for lo, hi in spans:
            if merged and lo < merged[-1][1]:
                merged[-1][1] = max(merged[-1][1], hi)
            else:
                merged.append([lo, hi])
        ok = 0
        with torch.cuda.device(self._rt.device):
            for lo, hi in merged:
                already = ctypes.c_int(0)
                rc = _cabi.load().xchg_host_register(ctypes.c_void_p(lo), hi - lo, ctypes.byref(already))
                if rc != 0:
                    self.failed += 1  # the copy still works: the driver stages pageable memory itself
                    continue
                ok += 1
                if not already.value:  # else: someone (an arena registered once) page-locked it before
                    self._ranges.append(lo)
                    self.registered_bytes += hi - lo
        return ok

    def release(self):
        if self._ranges:
            self._rt.synchronize()  # no copy may still be reading the pages
            with torch.cuda.device(self._rt.device):
                for lo in self._ranges:
                    _cabi.call("xchg_host_unregister", ctypes.c_void_p(lo))
        self._ranges = []
        self.registered_bytes = 0

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.release()
        return False


def stage_ipc(ex, buffers: Sequence, register: bool = False):
    """IPC streams (one per input block, as they sit in the object store) -> resident
    ``DeviceDataset``.  ``ex`` is a ``DeviceExchange``.  With ``register`` the wire buffers are
    page-locked for the duration of the copies (see the module docstring for when that pays)."""
    tables = [table_from_ipc(b) for b in buffers]
    if not register:
        return ex.stage(tables)
    with HostRegistration(ex.rt) as reg:
        reg.add(*buffers)
        ds = ex.stage(tables)
        ex.rt.synchronize()
    return ds


def export_ipc(ex, ds, memcopy_threads: int = 8) -> List[pa.Buffer]:
    """Resident dataset -> one IPC stream per output block, each in pinned host memory."""
    out = []
    for t in ex.export(ds):
        mock = pa.MockOutputStream()
        with pa.ipc.new_stream(mock, t.schema) as w:
            w.write_table(t)
        size = mock.size()
        host = ex.rt.host_empty(size)
        sink_buf = pa.py_buffer(host.numpy())
        sink = pa.FixedSizeBufferWriter(sink_buf)
        sink.set_memcopy_threads(max(1, int(memcopy_threads)))
        with pa.ipc.new_stream(sink, t.schema) as w:
            w.write_table(t)
        sink.close()
        out.append(sink_buf)
    return out


def sort_ipc(ex, buffers: Sequence, sort_key, num_outputs: Optional[int] = None,
             boundaries: Optional[list] = None, seed: int = 0, register: bool = False) -> List[pa.Buffer]:
    """``Dataset.sort`` from wire bytes to wire bytes: what a reducer-side drop-in hands back to
    the object store."""
    ds = stage_ipc(ex, buffers, register)
    return export_ipc(ex, ex.sort(ds, sort_key, num_outputs, boundaries, seed))

"""Wire format <-> device staging (cluster_anywhere_b200/wire.py): Arrow IPC
streams as the reference's plasma fallback writes them
(python/ray/_private/arrow_serialization.py:785-809) go in, sorted IPC streams come out;
compared with the oracle's exchange on the same blocks."""
import numpy as np
import pyarrow as pa
import pytest

from _golden import tables_equal
from cluster_anywhere_b200 import SortKey, wire
from cluster_anywhere_b200.exchange import DeviceExchange
from oracle import exchange_oracle as O


def _blocks(rng, sizes):
    out = []
    for n in sizes:
        s = ["w%03d" % int(i) * int(rng.integers(0, 3)) for i in rng.integers(0, 50, n)]
        t = pa.table({
            "k": pa.array(rng.integers(-30, 30, n), pa.int64()),
            "f": pa.array(rng.standard_normal(n).astype(np.float32), mask=rng.random(n) < 0.1),
            "s": pa.array(s, pa.string(), mask=rng.random(n) < 0.1),
            "p": pa.array([rng.bytes(12) for _ in range(n)], pa.binary()),
        })
        out.append(t.slice(3) if n > 10 else t)  # sliced on the way in: offsets do not start at 0
    return out


@pytest.mark.parametrize("register", [False, True])
def test_ipc_in_ipc_out_sort(engine, register):
    rng = np.random.default_rng(8)
    blocks = _blocks(rng, [700, 0, 1, 1300])
    wires = [wire.table_to_ipc(b) for b in blocks]
    # the reader hands out views of the wire bytes
    back = wire.table_from_ipc(wires[0])
    assert tables_equal(back, blocks[0])
    lo, hi = wires[0].address, wires[0].address + wires[0].size
    col = back["p"].chunk(0)
    assert lo <= col.buffers()[2].address < hi
    bounds = [(-10,), (0,), (0,), (17,)]
    ex = DeviceExchange()
    out = wire.sort_ipc(ex, wires, SortKey("k"), boundaries=bounds, register=register)
    got = [wire.table_from_ipc(b) for b in out]
    exp = O.sort_exchange(blocks, bounds, ["k"], [False])
    assert len(got) == len(exp)
    for g, e in zip(got, exp):
        assert tables_equal(g, e)


def test_ipc_file_format_and_plain_bytes(engine):
    t = pa.table({"k": pa.array([3, 1, 2], pa.int32())})
    sink = pa.BufferOutputStream()
    with pa.ipc.new_file(sink, t.schema) as w:
        w.write_table(t)
    as_file = sink.getvalue().to_pybytes()
    as_stream = bytearray(wire.table_to_ipc(t).to_pybytes())
    ex = DeviceExchange()
    for buf in (as_file, as_stream, memoryview(as_stream)):
        ds = wire.stage_ipc(ex, [buf])
        out = ex.export(ex.sort(ds, SortKey("k")))
        assert out[0]["k"].to_pylist() == [1, 2, 3]


@pytest.mark.gpu
def test_host_registration_makes_wire_bytes_pinned():
    import torch

    from cluster_anywhere_b200 import device as dev

    rt = dev.Runtime()
    a = np.arange(3 << 20, dtype=np.uint8)  # pageable
    b = np.arange(5000, dtype=np.uint8)
    with wire.HostRegistration(rt) as reg:
        assert reg.add(a, b, a[100:200]) >= 1
        assert reg.failed == 0 and reg.registered_bytes >= a.nbytes
        assert torch.from_numpy(a).is_pinned()
        d = torch.empty(a.size, dtype=torch.uint8, device=rt.device)
        d.copy_(torch.from_numpy(a), non_blocking=True)
        rt.synchronize()
        assert bool((d.cpu() == torch.from_numpy(a)).all())
        assert reg.add(a) == 1  # already registered: fine
    assert not torch.from_numpy(a).is_pinned()

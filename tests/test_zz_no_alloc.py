"""xchg_gather_offsets_static: the entry point for callers that cannot allocate.

Some hosts of this library only ever hold device buffers that another component
gave them (their GPU memory sits in someone else's pool, or they must not pay
cudaMalloc's device-wide stall inside their loop), so they cannot supply the
scratch workspace xchg_gather_offsets asks for.  The static variant keeps that
scratch in a fixed array inside the library.  Here it is called through the
C-ABI, twice in a row with different inputs (the second call finds the first
call's left-overs in the shared scratch), and must reproduce
xchg_gather_offsets bit for bit (integer offsets: exact equality, as everywhere
in this suite).  Kept in a file that sorts last so a `-x` run has already run
every other test; it also passes on its own in a fresh process.
"""
import ctypes

import numpy as np
import pytest


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


@pytest.mark.gpu
def test_static_scratch_variant_matches_workspace_variant():
    import torch

    from cluster_anywhere_b200 import _cabi

    _cabi._install_test_double(None)
    lib = _cabi.load()
    assert torch.cuda.is_available()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rng = np.random.default_rng(123)

    # two consecutive calls: a multi-tile gather, then a shorter one with other lengths
    for n_in, n_out, max_len in ((300_000, 1_000_003, 200), (5_000, 70_001, 17)):
        lens = rng.integers(0, max_len, n_in)
        in_off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)).to(dev)
        idx = torch.from_numpy(rng.integers(0, n_in, n_out).astype(np.int32)).to(dev)

        ws_bytes = ctypes.c_uint64()
        _cabi.check(lib.xchg_gather_offsets_workspace_bytes(n_out, ctypes.byref(ws_bytes)))
        ws = torch.empty(ws_bytes.value, dtype=torch.uint8, device=dev)
        ref = torch.full((n_out + 1,), -1, dtype=torch.int64, device=dev)
        _cabi.check(lib.xchg_gather_offsets(_vp(in_off), _vp(idx), n_out, _vp(ref), _vp(ws),
                                            ws_bytes.value, stream), "xchg_gather_offsets")

        got = torch.full((n_out + 1,), -2, dtype=torch.int64, device=dev)
        _cabi.check(lib.xchg_gather_offsets_static(_vp(in_off), _vp(idx), n_out, _vp(got), stream),
                    "xchg_gather_offsets_static")
        torch.cuda.synchronize()
        assert torch.equal(got, ref)
        # and both agree with the definition
        exp = np.concatenate([[0], np.cumsum(lens[idx.cpu().numpy()])])
        assert np.array_equal(got.cpu().numpy(), exp)

    # empty request, and a request beyond the built-in scratch is refused loudly
    got = torch.full((1,), 7, dtype=torch.int64, device=dev)
    _cabi.check(lib.xchg_gather_offsets_static(None, None, 0, _vp(got), stream))
    torch.cuda.synchronize()
    assert int(got.item()) == 0
    too_many = (64 << 20) + 1
    rc = lib.xchg_gather_offsets_static(_vp(got), _vp(got), too_many, _vp(got), stream)
    assert rc == -4  # XCHG_ERR_LIMIT
    assert b"XCHG_STATIC_GATHER_MAX_ROWS" in lib.xchg_last_error()


@pytest.mark.gpu
def test_static_scratch_variant_at_its_row_limit():
    """Exactly XCHG_STATIC_GATHER_MAX_ROWS rows: every one of the built-in scratch's tile
    slots is in use, and the last offset is still the sum of all gathered lengths."""
    import torch

    from cluster_anywhere_b200 import _cabi

    _cabi._install_test_double(None)
    lib = _cabi.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    n_in, n_out = 1000, 64 << 20
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    lens = torch.randint(0, 9, (n_in,), dtype=torch.int64, device=dev, generator=g)
    in_off = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(lens, 0)])
    idx = torch.randint(0, n_in, (n_out,), dtype=torch.int32, device=dev, generator=g)
    got = torch.empty(n_out + 1, dtype=torch.int64, device=dev)
    _cabi.check(lib.xchg_gather_offsets_static(_vp(in_off), _vp(idx), n_out, _vp(got), stream),
                "xchg_gather_offsets_static")
    torch.cuda.synchronize()
    exp = torch.cumsum(lens[idx.long()], 0)
    assert int(got[0].item()) == 0 and torch.equal(got[1:], exp)

"""The C-ABI library loads and exports every symbol include/xchg_b200.h declares
(no compute calls: this runs without a GPU)."""
import os
import re

from cluster_anywhere_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "xchg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(xchg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g

    g.build()
    lib = _cabi.load()
    declared = _declared()
    assert declared, "header declares nothing?"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/xchg_b200.h but not exported"
    # and the ctypes table binds exactly the declared set
    assert sorted(_cabi.exported_symbols()) == declared
    assert lib.xchg_abi_version() == 1


def test_no_cpu_fallback_without_device():
    """Product path must fail loudly when there is no CUDA device."""
    import pytest
    import torch

    from cluster_anywhere_b200 import XchgError, device

    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    device.reset_default_runtime()
    with pytest.raises(XchgError):
        device.default_runtime()

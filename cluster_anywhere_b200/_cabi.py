"""ctypes binding of libxchg_b200.so (the C-ABI declared in include/xchg_b200.h).

The library is built in-tree by ``__graft_entry__.build()``.  There is no CPU
fallback: if the shared object is missing, or a call fails, this module raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_int, c_int64, c_uint32, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libxchg_b200.so")

# key types (xchg_key_type)
I8, I16, I32, I64, U8, U16, U32, U64, F32, F64 = range(10)


class XchgError(RuntimeError):
    pass


_P = c_void_p


class MixedCol(ctypes.Structure):
    """xchg_mixed_col (include/xchg_b200.h): one key column of a tuple that mixes numeric
    and string columns."""
    _fields_ = [("kind", ctypes.c_int32), ("descending", ctypes.c_int32), ("norm", c_void_p),
                ("last_threshold", c_uint64), ("has_last", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("offsets", c_void_p), ("data", c_void_p), ("validity", c_void_p),
                ("bound_offsets", c_void_p), ("bound_data", c_void_p)]


_SIGNATURES = {
    "xchg_abi_version": (c_int, []),
    "xchg_last_error": (c_char_p, []),
    "xchg_device_count": (c_int, []),
    "xchg_launch_count": (c_uint64, []),
    "xchg_profile_enable": (c_int, [c_int]),
    "xchg_profile_read": (c_int, [c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_uint64)]),
    "xchg_sort_workspace_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_sort_column": (c_int, [_P, c_int, c_int, _P, c_int64, _P, c_uint64, _P, _P, _P, c_uint64, _P]),
    "xchg_normalize_keys": (c_int, [_P, c_int, c_int, _P, c_int64, _P, c_uint64, _P, _P]),
    "xchg_denormalize_keys": (c_int, [_P, c_int, c_int, c_uint64, _P, _P]),
    "xchg_set_l2_fetch_granularity": (c_int, [c_int]),
    "xchg_set_gather_ctas_per_sm": (c_int, [c_int]),
    "xchg_set_gather_stage_kb": (c_int, [c_int]),
    "xchg_set_sort_algorithm": (c_int, [c_int]),
    "xchg_partition_bounds": (c_int, [ctypes.POINTER(c_void_p), _P, _P, _P, c_int, c_uint64, _P, _P, _P, c_int, c_int, _P, _P]),
    "xchg_assign_partitions": (c_int, [ctypes.POINTER(c_void_p), _P, c_int, c_uint64, _P, _P, _P, c_int, c_int, _P, _P, _P]),
    "xchg_partition_bounds_mixed": (c_int, [ctypes.POINTER(MixedCol), c_int, c_uint64, _P, _P, _P, c_int, c_int, _P, _P]),
    "xchg_assign_partitions_mixed": (c_int, [ctypes.POINTER(MixedCol), c_int, c_uint64, _P, _P, _P, c_int, c_int, _P,
                                             _P, _P]),
    "xchg_sample_keys": (c_int, [_P, c_int, c_int, _P, c_int64, ctypes.POINTER(c_uint64), ctypes.POINTER(c_uint64),
                                 c_int, c_uint32, c_uint64, c_uint64, _P, c_uint64, ctypes.POINTER(c_uint64), _P]),
    "xchg_select_boundaries": (c_int, [_P, c_int, c_uint64, c_uint32, c_uint32, c_uint32, _P, _P, _P]),
    "xchg_partition_hist": (c_int, [_P, c_int, c_int, c_uint64, _P, ctypes.POINTER(c_uint32),
                                    ctypes.POINTER(ctypes.c_uint8), c_int, _P, _P]),
    "xchg_partition_workspace_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_partition_rows": (c_int, [_P, c_int, c_int, c_uint64, _P, ctypes.POINTER(c_uint32),
                                    ctypes.POINTER(ctypes.c_uint8), c_int, _P, _P, _P, _P, c_uint64, _P]),
    "xchg_copy_segments": (c_int, [c_int, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p),
                                   ctypes.POINTER(c_uint64), _P]),
    "xchg_signal_flags": (c_int, [ctypes.POINTER(c_void_p), c_int, c_uint32, _P]),
    "xchg_wait_flags": (c_int, [_P, c_int, c_uint32, c_uint64, _P, _P]),
    "xchg_merge_workspace_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_merge_sorted_runs": (c_int, [_P, _P, ctypes.POINTER(c_uint64), c_int, _P, _P, _P, c_uint64, _P, _P]),
    "xchg_shuffle_positions": (c_int, [c_uint64, c_uint64, c_uint64, c_uint64, c_int, _P, _P]),
    "xchg_invert_permutation": (c_int, [_P, c_uint64, c_uint64, _P, _P]),
    "xchg_even_split_dest": (c_int, [_P, c_uint64, c_uint64, c_uint32, _P, _P, _P]),
    "xchg_hash_table_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_hash_groups": (c_int, [_P, _P, c_int64, c_uint64, c_uint64, _P, c_uint64, _P, _P, _P,
                                 ctypes.POINTER(c_uint64), _P]),
    "xchg_group_ids": (c_int, [_P, _P, c_uint64, _P, c_uint64, c_uint64, _P, _P, _P]),
    "xchg_agg_stats_init": (c_int, [_P, _P, c_uint64, _P]),
    "xchg_key_minmax": (c_int, [_P, _P, c_int64, c_uint64, _P, ctypes.POINTER(c_uint64), _P]),
    "xchg_dense_groups_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_dense_groups": (c_int, [_P, _P, c_int64, c_uint64, c_uint64, c_uint64, _P, c_uint64, _P, _P, _P,
                                  ctypes.POINTER(c_uint64), _P]),
    "xchg_group_flags_fixed": (c_int, [_P, _P, c_int64, _P, c_uint64, _P, _P]),
    "xchg_group_flags_bytes": (c_int, [_P, _P, _P, c_int64, _P, c_uint64, _P, _P]),
    "xchg_group_first_rows": (c_int, [_P, c_uint64, c_uint64, _P, _P]),
    "xchg_groups_from_flags_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_groups_from_flags": (c_int, [_P, _P, c_uint64, _P, c_uint64, _P, _P, ctypes.POINTER(c_uint64), _P]),
    "xchg_agg_count_rows": (c_int, [_P, c_uint64, c_uint64, _P, _P]),
    "xchg_agg_accumulate": (c_int, [_P, _P, c_int, _P, c_int64, c_uint64, c_uint64, _P, c_int, _P, _P]),
    "xchg_agg_accumulate_m2": (c_int, [_P, _P, c_int, _P, c_int64, c_uint64, c_uint64, _P, _P, _P]),
    "xchg_agg_merge_rows": (c_int, [_P, _P, c_uint64, _P, _P]),
    "xchg_agg_merge_states": (c_int, [_P, _P, c_uint64, c_int, c_uint64, _P, _P, _P]),
    "xchg_agg_pack_states": (c_int, [c_uint64, _P, _P, _P, _P]),
    "xchg_agg_finalize": (c_int, [c_int, c_int, c_int, c_int, c_uint64, _P, _P, _P, _P, _P, _P]),
    "xchg_group_keys_out": (c_int, [_P, c_int, c_uint64, _P, _P]),
    "xchg_gather_bitmap_bytes": (c_int, [_P, c_int64, _P, c_uint64, _P, _P]),
    "xchg_bitmap_unpack": (c_int, [_P, c_int64, c_uint64, _P, _P]),
    "xchg_bitmap_pack": (c_int, [_P, c_uint64, _P, _P, _P]),
    "xchg_offsets_rebase": (c_int, [_P, c_uint64, c_int64, _P, _P]),
    "xchg_string_lengths": (c_int, [_P, c_uint64, _P, _P, _P]),
    "xchg_string_chunk_keys": (c_int, [_P, _P, c_uint64, c_uint32, _P, _P]),
    "xchg_string_tie_flags": (c_int, [_P, _P, _P, _P, _P, _P, c_uint64, c_uint64, _P, _P, _P]),
    "xchg_string_active_rows": (c_int, [_P, _P, _P, c_uint64, c_uint64, _P, _P, _P]),
    "xchg_string_run_heads": (c_int, [_P, _P, c_uint64, c_uint32, _P, _P]),
    "xchg_scatter_flagged": (c_int, [_P, _P, c_uint64, _P, _P]),
    "xchg_string_chunk_keys_rows": (c_int, [_P, _P, _P, c_uint64, c_uint32, _P, _P]),
    "xchg_scatter_u32": (c_int, [_P, _P, c_uint64, _P, _P]),
    "xchg_partition_bounds_bytes": (c_int, [_P, _P, _P, c_uint64, _P, _P, _P, c_int, c_int, _P, _P]),
    "xchg_assign_partitions_bytes": (c_int, [_P, _P, _P, c_uint64, _P, _P, _P, c_int, c_int, _P, _P, _P]),
    "xchg_gather_fixed": (c_int, [_P, c_uint32, _P, c_uint64, _P, _P]),
    "xchg_gather_bitmap": (c_int, [_P, c_int64, _P, c_uint64, _P, _P, _P]),
    "xchg_gather_offsets_workspace_bytes": (c_int, [c_uint64, ctypes.POINTER(c_uint64)]),
    "xchg_gather_offsets": (c_int, [_P, _P, c_uint64, _P, _P, c_uint64, _P]),
    "xchg_gather_offsets_static": (c_int, [_P, _P, c_uint64, _P, _P]),
    "xchg_gather_bytes": (c_int, [_P, _P, _P, c_uint64, _P, _P, _P]),
    "xchg_gather_spans": (c_int, [_P, _P, c_uint64, _P, _P, _P, c_uint64, _P]),
    "xchg_gather_bytes_spans": (c_int, [_P, _P, c_uint64, _P, _P, _P]),
    "xchg_gather_rows": (c_int, [_P, c_uint32, _P, c_uint64, _P, _P]),
    "xchg_gather_rows_multi": (c_int, [_P, c_uint32, _P, c_int, ctypes.POINTER(c_uint64), ctypes.POINTER(c_uint64),
                                       ctypes.POINTER(c_void_p), _P]),
    "xchg_host_register": (c_int, [_P, c_uint64, ctypes.POINTER(c_int)]),
    "xchg_host_unregister": (c_int, [_P]),
    "xchg_offsets_widen": (c_int, [_P, c_uint64, c_int64, _P, _P]),
    "xchg_offsets_narrow": (c_int, [_P, c_uint64, _P, _P]),
    "xchg_offsets_uniform32": (c_int, [c_uint64, c_uint32, _P, _P]),
    "xchg_offsets_uniform64": (c_int, [c_uint64, c_uint64, c_int64, _P, _P]),
    "xchg_offsets_check_uniform": (c_int, [_P, c_uint64, c_uint64, _P, _P]),
}

_lib = None
_test_double = None


def exported_symbols():
    """Names every build of the library must export (mirrors include/xchg_b200.h)."""
    return sorted(_SIGNATURES)


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _test_double is not None:
        return _test_double
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise XchgError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` -- this engine has no CPU fallback"
            )
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        if lib.xchg_abi_version() != 1:
            raise XchgError("libxchg_b200.so ABI version mismatch; rebuild")
        _lib = lib
    return _lib


def _install_test_double(obj):
    """tests/ only: route calls to a host-memory stand-in so that the Python host
    logic can be exercised without a GPU.  Never used by the product path."""
    global _test_double
    _test_double = obj


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().xchg_last_error()
        if isinstance(msg, bytes):
            msg = msg.decode("utf-8", "replace")
        raise XchgError(f"{what or 'xchg call'} failed (status {rc}): {msg}")


def call(name: str, *args):
    fn = getattr(load(), name)
    check(fn(*args), name)

"""Exchange operator surface (python/ray/data/_internal/planner/exchange/interfaces.py:14-102).

``ExchangeTaskSpec`` describes one exchange as a stateless ``map`` (one input
block -> ``output_num_blocks`` blocks + metadata LAST) and ``reduce`` (one block
from every mapper -> one output block + metadata); ``ExchangeTaskScheduler``
runs them.  Same names and calling conventions as the reference so that specs
and schedulers are interchangeable with the reference's.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

from ..block import BlockMetadata


class ExchangeTaskSpec:
    MAP_SUB_PROGRESS_BAR_NAME = "Shuffle Map"
    REDUCE_SUB_PROGRESS_BAR_NAME = "Shuffle Reduce"

    def __init__(self, map_args: Optional[List[Any]] = None, reduce_args: Optional[List[Any]] = None):
        self._map_args = map_args or []
        self._reduce_args = reduce_args or []
        assert isinstance(self._map_args, list)
        assert isinstance(self._reduce_args, list)

    @staticmethod
    def map(idx: int, block, output_num_blocks: int, *map_args) -> list:
        raise NotImplementedError

    @staticmethod
    def reduce(*reduce_args_and_mapper_outputs, partial_reduce: bool = False) -> Tuple[Any, BlockMetadata]:
        raise NotImplementedError


class ExchangeTaskScheduler:
    def __init__(self, exchange_spec: ExchangeTaskSpec):
        self._exchange_spec = exchange_spec

    def execute(self, refs: list, output_num_blocks: int, task_ctx=None,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None):
        raise NotImplementedError


# ---- RefBundle / TaskContext conventions shared by the schedulers ------------------------
# The reference's ``execute`` takes ``List[RefBundle]`` and a ``TaskContext`` and returns
# ``(List[RefBundle], StatsDict)`` (pull_based_shuffle_task_scheduler.py:31-150).  The
# schedulers here accept those, and -- for callers that hold plain blocks -- a plain list
# of blocks, in which case they return plain blocks.

def unwrap_refs(refs: list):
    """-> (blocks, input_owned, wrapped).  ``wrapped``: the input was RefBundles."""
    wrapped = any(hasattr(r, "block_refs") and hasattr(r, "owns_blocks") for r in refs)
    if not wrapped:
        return list(refs), True, False
    blocks = []
    for r in refs:
        blocks.extend(r.block_refs)
    return blocks, all(r.owns_blocks for r in refs), True


def wrap_outputs(blocks: list, metas: list, input_owned: bool, wrapped: bool) -> list:
    """One output bundle per reducer, in reducer order, owned iff every input was
    (pull_based_shuffle_task_scheduler.py:131-144)."""
    if not wrapped:
        return list(blocks)
    from ..block_accessor import B200ArrowBlockAccessor
    from ..planner import RefBundle

    out = []
    for b, m in zip(blocks, metas):
        if m is None or m.size_bytes is None:
            m = B200ArrowBlockAccessor.for_block(b).get_metadata(exec_stats=getattr(m, "exec_stats", None))
        out.append(RefBundle([(b, m)], owns_blocks=input_owned))
    return out


def sub_progress_bar(task_ctx, name: str):
    """The named sub progress bar of the task context; a context that carries bars must
    carry this one (asserted in the reference, pull_based...:76-79, push_based...:523-526)."""
    bars = getattr(task_ctx, "sub_progress_bar_dict", None)
    if bars is None:
        return None
    assert name in bars, bars
    return bars[name]

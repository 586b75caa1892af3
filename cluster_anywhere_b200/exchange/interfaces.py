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

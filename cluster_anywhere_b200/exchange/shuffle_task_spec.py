"""``ShuffleTaskSpec`` on the GPU
(python/ray/data/_internal/planner/exchange/shuffle_task_spec.py:15-132)."""
from __future__ import annotations

import logging
import math
from typing import Callable, Iterable, List, Optional, Tuple

import numpy as np
import pyarrow as pa

from ..block import BlockExecStats, BlockMetadata
from ..block_accessor import B200ArrowBlockAccessor
from ..block_builder import DelegatingBlockBuilder
from .interfaces import ExchangeTaskSpec

logger = logging.getLogger(__name__)

# python/ray/data/context.py:51
MAX_SAFE_BLOCK_SIZE_FACTOR = 1.5


class ShuffleTaskSpec(ExchangeTaskSpec):
    """Random shuffle (and plain repartition when ``random_shuffle=False``):
    map = permute the block and cut it into R equal slices, reduce = concatenate one
    slice per mapper and permute again."""

    SPLIT_REPARTITION_SUB_PROGRESS_BAR_NAME = "Split Repartition"

    def __init__(self, target_shuffle_max_block_size: int, random_shuffle: bool = False,
                 random_seed: Optional[int] = None,
                 upstream_map_fn: Optional[Callable[[Iterable], Iterable]] = None):
        super().__init__(
            map_args=[target_shuffle_max_block_size, upstream_map_fn, random_shuffle, random_seed],
            reduce_args=[random_shuffle, random_seed],
        )

    @staticmethod
    def map(idx: int, block, output_num_blocks: int, target_shuffle_max_block_size: int,
            upstream_map_fn, random_shuffle: bool, random_seed: Optional[int]) -> list:
        """shuffle_task_spec.py:41-104."""
        stats = BlockExecStats.builder()
        if upstream_map_fn:
            mapped = [B200ArrowBlockAccessor.for_block(b).to_arrow() for b in upstream_map_fn([block])]
            block = pa.concat_tables(mapped, promote_options="permissive") if mapped else pa.table({})
        acc = B200ArrowBlockAccessor.for_block(block)
        if acc.size_bytes() > MAX_SAFE_BLOCK_SIZE_FACTOR * target_shuffle_max_block_size:
            logger.warning(
                "Input block to map task has size %dMiB, which exceeds "
                "target_shuffle_max_block_size=%dMiB.", acc.size_bytes() // (1024 * 1024),
                target_shuffle_max_block_size // (1024 * 1024))
        if random_shuffle:
            seed_i = random_seed + idx if random_seed is not None else None
            acc = B200ArrowBlockAccessor.for_block(acc.random_shuffle(seed_i))
        slice_sz = max(1, math.ceil(acc.num_rows() / output_num_blocks))
        slices = [acc.slice(i * slice_sz, (i + 1) * slice_sz) for i in range(output_num_blocks)]
        if random_shuffle:
            # the ORDER of the R slices is shuffled on the host exactly as the reference
            # does (a list of R table handles, :97-99) -- so reducer j receives the
            # same slice sizes as under the reference
            np.random.RandomState(seed_i).shuffle(slices)
        num_rows = sum(B200ArrowBlockAccessor.for_block(s).num_rows() for s in slices)
        assert num_rows == acc.num_rows(), (num_rows, acc.num_rows())
        return slices + [acc.get_metadata(exec_stats=stats.build())]

    @staticmethod
    def reduce(random_shuffle: bool, random_seed: Optional[int], *mapper_outputs,
               partial_reduce: bool = False) -> Tuple[pa.Table, BlockMetadata]:
        """shuffle_task_spec.py:106-132."""
        stats = BlockExecStats.builder()
        builder = DelegatingBlockBuilder()
        for b in mapper_outputs:
            b = B200ArrowBlockAccessor.for_block(b).to_arrow()
            if b.num_columns > 0:
                builder.add_block(b)
        new_block = builder.build()
        acc = B200ArrowBlockAccessor.for_block(new_block)
        if random_shuffle:
            new_block = acc.random_shuffle(random_seed if random_seed is not None else None)
            acc = B200ArrowBlockAccessor.for_block(new_block)
        meta = BlockMetadata(acc.num_rows(), acc.size_bytes(), acc.schema(), None, stats.build())
        from ..block_accessor import want_pandas_output

        if want_pandas_output(mapper_outputs):  # DataFrame blocks in, a DataFrame out
            new_block = B200ArrowBlockAccessor.for_block(new_block).to_arrow().to_pandas()
        return new_block, meta

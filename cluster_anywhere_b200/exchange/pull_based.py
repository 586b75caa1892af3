"""Pull-based schedule (python/ray/data/_internal/planner/exchange/
pull_based_shuffle_task_scheduler.py:31-150): every map task, then reducer j
receives output j of every mapper.  The reference submits Ray tasks; on one GPU
the tasks run back to back in this process (kernels of consecutive tasks overlap
on the stream)."""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

from .interfaces import (ExchangeTaskScheduler, ExchangeTaskSpec, sub_progress_bar, unwrap_refs,
                         wrap_outputs)


class PullBasedShuffleTaskScheduler(ExchangeTaskScheduler):
    def execute(self, refs: list, output_num_blocks: int, task_ctx=None,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None,
                _debug_limit_execution_to_num_blocks: Optional[int] = None):
        blocks, input_owned, wrapped = unwrap_refs(refs)
        input_num_blocks = len(blocks)
        spec = self._exchange_spec
        map_bar = sub_progress_bar(task_ctx, ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME)
        reduce_bar = sub_progress_bar(task_ctx, ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME)
        if _debug_limit_execution_to_num_blocks is not None:
            blocks = blocks[:_debug_limit_execution_to_num_blocks]
        map_out, map_meta = [], []
        for i, b in enumerate(blocks):
            res = spec.map(i, b, output_num_blocks, *spec._map_args)
            map_meta.append(res[-1])
            map_out.append(res[:-1])
        if _debug_limit_execution_to_num_blocks is not None:
            while len(map_out) < input_num_blocks:  # repeat the first map task's results (:100-103)
                map_out.append(list(map_out[0]))
            output_num_blocks = _debug_limit_execution_to_num_blocks
        if map_bar is not None:
            map_meta = map_bar.fetch_until_complete(map_meta)
        out_blocks, out_meta = [], []
        for j in range(output_num_blocks):
            blk, meta = spec.reduce(*spec._reduce_args, *[m[j] for m in map_out])
            out_blocks.append(blk)
            out_meta.append(meta)
        if reduce_bar is not None:
            out_meta = reduce_bar.fetch_until_complete(out_meta)
        stats = {"map": map_meta, "reduce": out_meta}
        return wrap_outputs(out_blocks, out_meta, input_owned, wrapped), stats

"""Push-based (Exoshuffle) schedule
(python/ray/data/_internal/planner/exchange/push_based_shuffle_task_scheduler.py:28-828).

The reference bounds the number of intermediate objects of a large shuffle by running
the map tasks in ROUNDS: each round's map outputs are immediately combined by a few
merge tasks (``reduce(..., partial_reduce=True)``), one merge task serving a contiguous
run of reducers, and only the per-round merge results reach the final reduce tasks.
Round r+1's maps are pipelined with round r's merges.

Everything that decides WHAT is computed is restated here so that results, stats keys
and the schedule's arithmetic match the reference:

* ``MergeSchedule``        which reducers a merge task serves (:28-109)
* ``compute_push_schedule`` merge-task count / placement, maps per round, rounds (:733-815)
* ``PushBasedShuffleTaskScheduler.execute``  rounds of map + merge, then reduce; outputs
  ordered by reducer index; stats ``{"map", "merge", "reduce"}`` (:435-653)

What differs is WHERE tasks run: the reference submits Ray tasks with node-affinity
hints; on one GPU box the tasks of a round run back to back in this process (kernels of
consecutive tasks overlap on the stream), so merge placement is bookkeeping only --
NVSwitch makes every GPU equally close, which is the case the reference's placement
logic degenerates to on a single node.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, Iterator, List, Optional

from ..block import BlockExecStats, BlockMetadata
from ..block_accessor import B200ArrowBlockAccessor
from .interfaces import (ExchangeTaskScheduler, ExchangeTaskSpec, sub_progress_bar, unwrap_refs,
                         wrap_outputs)


class MergeSchedule:
    """Assignment of the ``output_num_blocks`` reducers to the merge tasks of a round:
    contiguous runs, the first ``extra`` merge tasks serving one reducer more."""

    def __init__(self, output_num_blocks: int, num_merge_tasks_per_round: int):
        self.output_num_blocks = output_num_blocks
        per, extra = divmod(output_num_blocks, num_merge_tasks_per_round)
        if per == 0:
            # fewer reducers than merge tasks: one reducer per merge task, drop the rest
            num_merge_tasks_per_round, per, extra = extra, 1, 0
        self.num_merge_tasks_per_round = num_merge_tasks_per_round
        self.num_reducers_per_merger = per
        self._extra = extra

    def get_num_reducers_per_merge_idx(self, merge_idx: int) -> int:
        assert merge_idx < self.num_merge_tasks_per_round
        return self.num_reducers_per_merger + (1 if merge_idx < self._extra else 0)

    def first_reducer_of(self, merge_idx: int) -> int:
        return merge_idx * self.num_reducers_per_merger + min(merge_idx, self._extra)

    def get_merge_idx_for_reducer_idx(self, reducer_idx: int) -> int:
        wide = (self.num_reducers_per_merger + 1) * self._extra  # reducers under the wider merge tasks
        if reducer_idx < wide:
            merge_idx = reducer_idx // (self.num_reducers_per_merger + 1)
        else:
            merge_idx = self._extra + (reducer_idx - wide) // self.num_reducers_per_merger
        assert merge_idx < self.num_merge_tasks_per_round
        return merge_idx

    def round_robin_reduce_idx_iterator(self) -> Iterator[int]:
        """Reducer indices in the order the reduce tasks are submitted: the k-th
        reducer of every merge task, for k = 0, 1, ... (spreads load over mergers)."""
        emitted, k = 0, 0
        while emitted < self.output_num_blocks:
            for m in range(self.num_merge_tasks_per_round):
                if k < self.get_num_reducers_per_merge_idx(m):
                    yield self.first_reducer_of(m) + k
                    emitted += 1
            k += 1


@dataclass
class PushStage:
    num_rounds: int
    num_map_tasks_per_round: int
    merge_task_placement: List[str]
    merge_schedule: MergeSchedule

    def get_estimated_num_refs(self) -> int:
        """Objects alive in the reference's object store for this schedule (:144-156)."""
        m = self.merge_schedule.num_merge_tasks_per_round
        inputs = self.num_rounds * self.num_map_tasks_per_round
        return inputs + inputs * m + self.merge_schedule.output_num_blocks


def compute_push_schedule(num_cpus_per_node_map: Dict[str, int], num_input_blocks: int,
                          merge_factor: float, num_output_blocks: int) -> PushStage:
    """How many merge tasks run per round and where, how many map tasks share a round,
    and how many rounds there are (push_based_shuffle_task_scheduler.py:733-815): every
    merge task is grouped with ``merge_factor`` map tasks, groups are dealt to nodes by
    their task slots, left-over slots of several nodes may form one more group, and the
    slots not used by merge tasks run map tasks."""
    slots_total = sum(num_cpus_per_node_map.values())
    group = merge_factor + 1
    merge_tasks_wanted = math.ceil(num_input_blocks / merge_factor)
    placement: List[str] = []
    leftover = 0
    inputs_left = num_input_blocks
    for node, slots in num_cpus_per_node_map.items():
        usable = min(slots, inputs_left)
        here = min(round(usable / group), merge_tasks_wanted)
        placement += [node] * here
        inputs_left -= merge_factor * here
        leftover += slots - group * here
        if leftover >= group and len(placement) < merge_tasks_wanted:
            placement.append(node)
            leftover -= group
            inputs_left -= merge_factor
        inputs_left = max(0, inputs_left)
    if not placement:  # tiny dataset: still one merge task
        for node, slots in num_cpus_per_node_map.items():
            if slots >= 1:
                placement.append(node)
                break
    assert placement, "no task slots available"
    maps_per_round = max(1, min(slots_total - len(placement), num_input_blocks))
    rounds = math.ceil(num_input_blocks / maps_per_round)
    return PushStage(rounds, maps_per_round, placement, MergeSchedule(num_output_blocks, len(placement)))


def _num_rows(block) -> int:
    return B200ArrowBlockAccessor.for_block(block).num_rows()


class PushBasedShuffleTaskScheduler(ExchangeTaskScheduler):
    """Rounds of map + merge tasks, then the reduce tasks (:400-653)."""

    def execute(self, refs: list, output_num_blocks: int, task_ctx=None,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None,
                merge_factor: float = 2,
                _debug_limit_execution_to_num_blocks: Optional[int] = None,
                num_cpus_per_node_map: Optional[Dict[str, int]] = None):
        import os

        spec = self._exchange_spec
        blocks, input_owned, wrapped = unwrap_refs(refs)
        map_bar = sub_progress_bar(task_ctx, ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME)
        reduce_bar = sub_progress_bar(task_ctx, ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME)
        if _debug_limit_execution_to_num_blocks is not None:
            blocks = blocks[:_debug_limit_execution_to_num_blocks]
        if num_cpus_per_node_map is None:
            # one box: its task slots (the reference reads ray.state, :818-828)
            num_cpus_per_node_map = {"node0": max(2, len(os.sched_getaffinity(0)))}
        stage = compute_push_schedule(num_cpus_per_node_map, len(blocks), merge_factor, output_num_blocks)
        sched = stage.merge_schedule
        n_merge = sched.num_merge_tasks_per_round

        map_meta: List[BlockMetadata] = []
        merge_meta: List[BlockMetadata] = []
        # merge_results[merge_idx][round] = list of partially reduced blocks, one per
        # reducer served by that merge task
        merge_results: List[List[List[Any]]] = [[] for _ in range(n_merge)]
        for rnd in range(stage.num_rounds):
            lo = rnd * stage.num_map_tasks_per_round
            round_blocks = blocks[lo:lo + stage.num_map_tasks_per_round]
            if not round_blocks:
                break
            # ---- map tasks of this round: R outputs each, dealt to the merge tasks ----
            per_merge_inputs: List[List[List[Any]]] = [[] for _ in range(n_merge)]
            for i, blk in enumerate(round_blocks):
                outs = spec.map(lo + i, blk, output_num_blocks, *spec._map_args)
                map_meta.append(outs[-1])
                outs = outs[:-1]
                assert len(outs) == output_num_blocks, (len(outs), output_num_blocks)
                for m in range(n_merge):
                    first = sched.first_reducer_of(m)
                    per_merge_inputs[m].append(outs[first:first + sched.get_num_reducers_per_merge_idx(m)])
            # ---- merge tasks: partial reduce of every served reducer's column (:696-730) ----
            for m in range(n_merge):
                stats = BlockExecStats.builder()
                merged, rows = [], 0
                for column in zip(*per_merge_inputs[m]):
                    blk, _ = spec.reduce(*spec._reduce_args, *column, partial_reduce=True)
                    merged.append(blk)
                    rows += _num_rows(blk)
                merge_results[m].append(merged)
                merge_meta.append(BlockMetadata(rows, None, None, None, stats.build()))
        # ---- reduce tasks, submitted round-robin over the mergers, returned by index ----
        results: Dict[int, Any] = {}
        reduce_meta: Dict[int, BlockMetadata] = {}
        limit = output_num_blocks
        if _debug_limit_execution_to_num_blocks is not None:
            limit = min(limit, _debug_limit_execution_to_num_blocks)
        for j in sched.round_robin_reduce_idx_iterator():
            if j >= limit:
                continue
            m = sched.get_merge_idx_for_reducer_idx(j)
            k = j - sched.first_reducer_of(m)
            inputs = [per_round[k] for per_round in merge_results[m]]
            blk, meta = spec.reduce(*spec._reduce_args, *inputs)
            results[j], reduce_meta[j] = blk, meta
        order = sorted(results)
        if map_bar is not None:
            map_bar.update(len(map_meta))
        metas = [reduce_meta[j] for j in order]
        if reduce_bar is not None:
            metas = reduce_bar.fetch_until_complete(metas)
        stats = {"map": map_meta, "merge": merge_meta, "reduce": metas}
        return wrap_outputs([results[j] for j in order], metas, input_owned, wrapped), stats

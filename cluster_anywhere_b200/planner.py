"""Planner / operator surface of the all-to-all exchange.

Mirrors, with the reference's names, argument order and return shapes:

* ``DataContext`` / ``ShuffleStrategy``            python/ray/data/context.py:27-32, 480-487
* ``RefBundle``                                    _internal/execution/interfaces/ref_bundle.py:13-128
* ``TaskContext``                                  _internal/execution/interfaces/task_context.py:11-44
* logical operators ``Sort`` / ``RandomShuffle`` / ``Aggregate`` / ``Repartition`` /
  ``RandomizeBlocks``                              _internal/logical/operators/all_to_all_operator.py
* ``generate_sort_fn``                             _internal/planner/sort.py:21-85
* ``generate_random_shuffle_fn``                   _internal/planner/random_shuffle.py:20-83
* ``generate_aggregate_fn``                        _internal/planner/aggregate.py:24-98
* ``generate_repartition_fn``                      _internal/planner/repartition.py:23-83
* ``generate_randomize_blocks_fn``                 _internal/planner/randomize_blocks.py:12-51
* ``plan_all_to_all_op`` / ``AllToAllOperator``    _internal/planner/plan_all_to_all_op.py:23-105,
                                                   _internal/execution/operators/base_physical_operator.py:42-153

What differs is only where blocks live and where tasks run: a "block reference" here is
the block itself (there is no object store), tasks run in this process on the GPU, and a
third shuffle strategy, ``ShuffleStrategy.B200_DEVICE_EXCHANGE``, hands the whole exchange
to ``DeviceExchange`` (one staging of all blocks, device-resident map / exchange / reduce,
all GPUs of the process group) instead of running the map and reduce tasks one by one.
"""
from __future__ import annotations

import enum
import random
from dataclasses import dataclass, field
from functools import partial
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import pyarrow as pa

from .block import BlockMetadata, empty_table
from .block_accessor import B200ArrowBlockAccessor
from .sort_key import SortKey

StatsDict = Dict[str, list]

# python/ray/data/context.py:53-64
DEFAULT_TARGET_MAX_BLOCK_SIZE = 128 * 1024 * 1024
DEFAULT_SHUFFLE_TARGET_MAX_BLOCK_SIZE = 1024 * 1024 * 1024


class ShuffleStrategy(str, enum.Enum):
    """python/ray/data/context.py:27-32, plus the device-resident strategy."""

    SORT_SHUFFLE_PULL_BASED = "sort_shuffle_pull_based"
    SORT_SHUFFLE_PUSH_BASED = "sort_shuffle_push_based"
    HASH_SHUFFLE = "hash_shuffle"
    B200_DEVICE_EXCHANGE = "b200_device_exchange"


class DataContext:
    """The fields of ``ray.data.DataContext`` this path reads."""

    _current: Optional["DataContext"] = None

    def __init__(self, shuffle_strategy: ShuffleStrategy = ShuffleStrategy.SORT_SHUFFLE_PULL_BASED,
                 target_max_block_size: int = DEFAULT_TARGET_MAX_BLOCK_SIZE,
                 target_shuffle_max_block_size: int = DEFAULT_SHUFFLE_TARGET_MAX_BLOCK_SIZE,
                 eager_free: bool = True):
        self.shuffle_strategy = shuffle_strategy
        self.target_max_block_size = target_max_block_size
        self.target_shuffle_max_block_size = target_shuffle_max_block_size
        self.eager_free = eager_free
        self._kv_configs: Dict[str, Any] = {}

    @staticmethod
    def get_current() -> "DataContext":
        if DataContext._current is None:
            DataContext._current = DataContext()
        return DataContext._current

    @staticmethod
    def _set_current(ctx: "DataContext") -> None:
        DataContext._current = ctx

    # context.py:434-452: the deprecated boolean view of the strategy
    @property
    def use_push_based_shuffle(self) -> bool:
        return self.shuffle_strategy == ShuffleStrategy.SORT_SHUFFLE_PUSH_BASED

    @use_push_based_shuffle.setter
    def use_push_based_shuffle(self, value: bool) -> None:
        self.shuffle_strategy = (ShuffleStrategy.SORT_SHUFFLE_PUSH_BASED if value
                                 else ShuffleStrategy.SORT_SHUFFLE_PULL_BASED)

    def get_config(self, key: str, default: Any = None) -> Any:
        return self._kv_configs.get(key, default)

    def set_config(self, key: str, value: Any) -> None:
        self._kv_configs[key] = value

    def remove_config(self, key: str) -> None:
        self._kv_configs.pop(key, None)


@dataclass
class RefBundle:
    """A group of blocks and their metadata (ref_bundle.py:13-128).  ``blocks`` holds
    ``(block, BlockMetadata)`` pairs; ``blocks`` and ``owns_blocks`` cannot be reassigned."""

    blocks: Tuple[Tuple[Any, BlockMetadata], ...]
    owns_blocks: bool
    output_split_idx: Optional[int] = None

    def __post_init__(self):
        if not isinstance(self.blocks, tuple):
            object.__setattr__(self, "blocks", tuple(self.blocks))
        for b in self.blocks:
            assert isinstance(b, tuple), b
            assert len(b) == 2, b
            assert isinstance(b[1], BlockMetadata), b
            if b[1].size_bytes is None:
                raise ValueError("The size in bytes of the block must be known: {}".format(b))

    def __setattr__(self, key, value):
        if hasattr(self, key) and key in ["blocks", "owns_blocks"]:
            raise ValueError(f"The `{key}` field of RefBundle cannot be updated.")
        object.__setattr__(self, key, value)

    @property
    def block_refs(self) -> List[Any]:
        return [b for b, _ in self.blocks]

    @property
    def metadata(self) -> List[BlockMetadata]:
        return [m for _, m in self.blocks]

    def num_rows(self) -> Optional[int]:
        total = 0
        for m in self.metadata:
            if m.num_rows is None:
                return None
            total += m.num_rows
        return total

    def size_bytes(self) -> int:
        return sum(m.size_bytes for m in self.metadata)

    def destroy_if_owned(self) -> int:
        """Bytes freed: the blocks are plain host tables here, so dropping the bundle is
        the release; the accounting follows the reference (:96-108)."""
        should_free = self.owns_blocks and DataContext.get_current().eager_free
        return self.size_bytes() if should_free else 0

    @staticmethod
    def from_block(block, owns_blocks: bool = True) -> "RefBundle":
        acc = B200ArrowBlockAccessor.for_block(block)
        return RefBundle([(block, acc.get_metadata())], owns_blocks=owns_blocks)


class ProgressBar:
    """The slice of _internal/progress_bar.py the exchange touches: schedulers look the
    bars up by name in ``TaskContext.sub_progress_bar_dict``, wait on them and update them."""

    def __init__(self, name: str, total: int = 1, unit: str = "row", position: int = 0):
        self.name, self.total, self.unit, self.position = name, total, unit, position
        self.progress = 0
        self.description = name
        self.closed = False

    def set_description(self, name: str) -> None:
        self.description = name

    def update(self, i: int = 0, total: Optional[int] = None) -> None:
        self.progress += i
        if total is not None:
            self.total = total

    def fetch_until_complete(self, refs: list) -> list:
        self.update(len(refs))
        return list(refs)

    def block_until_complete(self, remaining: list) -> None:
        self.update(len(remaining))

    def close(self) -> None:
        self.closed = True


@dataclass
class TaskContext:
    """task_context.py:11-44."""

    task_idx: int
    sub_progress_bar_dict: Optional[Dict[str, ProgressBar]] = None
    upstream_map_transformer: Optional[Any] = None
    upstream_map_ray_remote_args: Optional[Dict[str, Any]] = None
    target_max_block_size: Optional[int] = None
    kwargs: Dict[str, Any] = field(default_factory=dict)


def unify_block_metadata_schema(metadata: Sequence[BlockMetadata]) -> Optional[pa.Schema]:
    """_internal/util.py:unify_block_metadata_schema: one schema for all non-empty blocks
    (Arrow schemas are unified with type promotion)."""
    schemas = [m.schema for m in metadata
               if m.schema is not None and (m.num_rows is None or m.num_rows > 0)]
    schemas = [s for s in schemas if isinstance(s, pa.Schema) and len(s.names) > 0]
    if not schemas:
        return None
    return pa.unify_schemas(schemas, promote_options="permissive")


def to_stats(metas: Sequence[BlockMetadata]) -> list:
    """StatsDict values are lists of per-task BlockMetadata (stats.py:108-120 turns their
    exec_stats into the ``SortMap`` / ``SortReduce`` ... rows of ``ds.stats()``)."""
    return list(metas)


# ---- logical operators (all_to_all_operator.py) ---------------------------------------

class LogicalOperator:
    def __init__(self, name: str, input_dependencies: List["LogicalOperator"], num_outputs: Optional[int] = None):
        self._name = name
        self._input_dependencies = input_dependencies
        self._num_outputs = num_outputs

    @property
    def name(self) -> str:
        return self._name

    @property
    def input_dependencies(self) -> List["LogicalOperator"]:
        return self._input_dependencies


class InputData(LogicalOperator):
    """Leaf: blocks already in memory (logical/operators/input_data_operator.py)."""

    def __init__(self, input_data: Optional[List[RefBundle]] = None):
        super().__init__("InputData", [], len(input_data) if input_data is not None else None)
        self.input_data = input_data


class AbstractAllToAll(LogicalOperator):
    def __init__(self, name: str, input_op: LogicalOperator, num_outputs: Optional[int] = None,
                 sub_progress_bar_names: Optional[List[str]] = None,
                 ray_remote_args: Optional[Dict[str, Any]] = None):
        super().__init__(name, [input_op], num_outputs)
        self._num_outputs = num_outputs
        self._ray_remote_args = ray_remote_args or {}
        self._sub_progress_bar_names = sub_progress_bar_names


def _exchange_bar_names() -> List[str]:
    from .exchange.interfaces import ExchangeTaskSpec

    return [ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME, ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME]


class RandomizeBlocks(AbstractAllToAll):
    def __init__(self, input_op: LogicalOperator, seed: Optional[int] = None):
        super().__init__("RandomizeBlockOrder", input_op)
        self._seed = seed


class RandomShuffle(AbstractAllToAll):
    def __init__(self, input_op: LogicalOperator, name: str = "RandomShuffle", seed: Optional[int] = None,
                 ray_remote_args: Optional[Dict[str, Any]] = None):
        super().__init__(name, input_op, sub_progress_bar_names=_exchange_bar_names(),
                         ray_remote_args=ray_remote_args)
        self._seed = seed


class Repartition(AbstractAllToAll):
    def __init__(self, input_op: LogicalOperator, num_outputs: int, shuffle: bool):
        from .exchange.shuffle_task_spec import ShuffleTaskSpec

        names = (_exchange_bar_names() if shuffle
                 else [ShuffleTaskSpec.SPLIT_REPARTITION_SUB_PROGRESS_BAR_NAME])
        super().__init__("Repartition", input_op, num_outputs=num_outputs, sub_progress_bar_names=names)
        self._shuffle = shuffle


class Sort(AbstractAllToAll):
    def __init__(self, input_op: LogicalOperator, sort_key: SortKey, batch_format: Optional[str] = "default"):
        from .exchange.sort_task_spec import SortTaskSpec

        super().__init__("Sort", input_op,
                         sub_progress_bar_names=[SortTaskSpec.SORT_SAMPLE_SUB_PROGRESS_BAR_NAME]
                         + _exchange_bar_names())
        self._sort_key = sort_key
        self._batch_format = batch_format


class Aggregate(AbstractAllToAll):
    def __init__(self, input_op: LogicalOperator, key: Optional[str], aggs: list,
                 num_partitions: Optional[int] = None, batch_format: Optional[str] = "default"):
        # num_partitions is carried like the reference carries it (logical/operators/
        # all_to_all_operator.py:140-165); its planner does not read it either
        from .exchange.sort_task_spec import SortTaskSpec

        super().__init__("Aggregate", input_op,
                         sub_progress_bar_names=[SortTaskSpec.SORT_SAMPLE_SUB_PROGRESS_BAR_NAME]
                         + _exchange_bar_names())
        self._key = key
        self._aggs = aggs
        self._num_partitions = num_partitions
        self._batch_format = batch_format


# ---- AllToAllTransformFn factories -------------------------------------------------------

AllToAllTransformFn = Callable[[List[RefBundle], TaskContext], Tuple[List[RefBundle], StatsDict]]


def _flatten(refs: List[RefBundle]):
    blocks, metadata = [], []
    for ref_bundle in refs:
        blocks.extend(ref_bundle.block_refs)
        metadata.extend(ref_bundle.metadata)
    return blocks, metadata


def _device_world(group=None) -> int:
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size(group)
    except Exception:
        pass
    return 1


def _bundles_of(out_blocks: List[pa.Table], input_owned: bool, stats_builder) -> Tuple[List[RefBundle], list]:
    metas = []
    exec_stats = stats_builder.build()
    for b in out_blocks:
        metas.append(B200ArrowBlockAccessor.for_block(b).get_metadata(exec_stats=exec_stats))
    return [RefBundle([(b, m)], owns_blocks=input_owned) for b, m in zip(out_blocks, metas)], metas


def _scheduler_for(data_context: DataContext, spec):
    from .exchange.pull_based import PullBasedShuffleTaskScheduler
    from .exchange.push_based import PushBasedShuffleTaskScheduler

    if data_context.shuffle_strategy == ShuffleStrategy.SORT_SHUFFLE_PUSH_BASED:
        return PushBasedShuffleTaskScheduler(spec)
    return PullBasedShuffleTaskScheduler(spec)


def generate_sort_fn(sort_key: SortKey, batch_format: str, data_context: DataContext,
                     _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None) -> AllToAllTransformFn:
    """planner/sort.py:21-85.  R = number of input blocks; user boundaries are wrapped into
    1-tuples and reversed for a descending primary key (:55-62)."""
    from .block import BlockExecStats
    from .exchange.sort_task_spec import SortTaskSpec

    def fn(sort_key: SortKey, refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        blocks, metadata = _flatten(refs)
        device = data_context.shuffle_strategy == ShuffleStrategy.B200_DEVICE_EXCHANGE
        # (one process per GPU: a rank that holds no block still takes part in the collectives)
        collective = device and _device_world() > 1
        if len(blocks) == 0 and not collective:
            return (blocks, {})
        schema = unify_block_metadata_schema(metadata)
        if schema is not None or not collective:
            sort_key.validate_schema(schema)
        num_mappers = len(blocks)
        num_outputs = num_mappers
        boundaries = None
        if sort_key.boundaries:
            boundaries = [(b,) for b in sort_key.boundaries]
            if sort_key.get_descending()[0]:
                boundaries = boundaries[::-1]
            num_outputs = len(boundaries) + 1
        if device:
            # whole exchange on the device(s): boundaries are sampled there (K1)
            from .exchange.device_exchange import DeviceExchange

            st = BlockExecStats.builder()
            ex = DeviceExchange()
            seed = int(ctx.kwargs.get("sample_seed", 0))
            if _device_world() > 1:
                out = ex.sort_blocks_distributed(blocks, sort_key, boundaries=boundaries, seed=seed)
            else:
                out = ex.sort_blocks(blocks, sort_key, num_outputs=num_outputs, boundaries=boundaries, seed=seed)
            input_owned = all(b.owns_blocks for b in refs)
            for name in (ctx.sub_progress_bar_dict or {}):
                ctx.sub_progress_bar_dict[name].update(len(out))
            bundles, metas = _bundles_of(out, input_owned, st)
            return bundles, {"map": to_stats(metadata), "reduce": to_stats(metas)}
        if boundaries is None:
            sample_bar = (ctx.sub_progress_bar_dict or {}).get(SortTaskSpec.SORT_SAMPLE_SUB_PROGRESS_BAR_NAME)
            if ctx.sub_progress_bar_dict is not None:
                assert sample_bar is not None, ctx.sub_progress_bar_dict
            boundaries = SortTaskSpec.sample_boundaries(blocks, sort_key, num_outputs, sample_bar)
        sort_spec = SortTaskSpec(boundaries=boundaries, sort_key=sort_key, batch_format=batch_format)
        scheduler = _scheduler_for(data_context, sort_spec)
        return scheduler.execute(
            refs, num_outputs, ctx,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    return partial(fn, sort_key)


def generate_random_shuffle_fn(data_context: DataContext, seed: Optional[int],
                               num_outputs: Optional[int] = None,
                               ray_remote_args: Optional[Dict[str, Any]] = None,
                               _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None
                               ) -> AllToAllTransformFn:
    """planner/random_shuffle.py:20-83."""
    from .block import BlockExecStats
    from .exchange.shuffle_task_spec import ShuffleTaskSpec

    def fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        num_input_blocks = sum(len(r.blocks) for r in refs)
        map_transformer = ctx.upstream_map_transformer
        upstream_map_fn = None
        remote_args = ray_remote_args
        if map_transformer:
            # the fused upstream map must not slice its output: the shuffle fuses it back
            map_transformer.set_target_max_block_size(float("inf"))

            def upstream_map_fn(blocks):
                return map_transformer.apply_transform(blocks, ctx)

            remote_args = ctx.upstream_map_ray_remote_args
        if data_context.shuffle_strategy == ShuffleStrategy.B200_DEVICE_EXCHANGE:
            from .exchange.device_exchange import DeviceExchange

            st = BlockExecStats.builder()
            blocks, metadata = _flatten(refs)
            if upstream_map_fn is not None:
                blocks = [B200ArrowBlockAccessor.for_block(b).to_arrow() for blk in blocks
                          for b in upstream_map_fn([blk])]
            if len(blocks) == 0 and _device_world() <= 1:
                return ([], {})
            out = DeviceExchange().random_shuffle_blocks(blocks, seed, num_outputs)
            bundles, metas = _bundles_of(out, all(b.owns_blocks for b in refs), st)
            return bundles, {"map": to_stats(metadata), "reduce": to_stats(metas)}
        shuffle_spec = ShuffleTaskSpec(ctx.target_max_block_size, random_shuffle=True, random_seed=seed,
                                       upstream_map_fn=upstream_map_fn)
        if data_context.shuffle_strategy == ShuffleStrategy.SORT_SHUFFLE_PUSH_BASED:
            if num_outputs is not None:
                raise NotImplementedError("Push-based shuffle doesn't support setting num_blocks yet.")
        scheduler = _scheduler_for(data_context, shuffle_spec)
        return scheduler.execute(
            refs, num_outputs or num_input_blocks, task_ctx=ctx,
            map_ray_remote_args=remote_args, reduce_ray_remote_args=remote_args,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    return fn


def generate_aggregate_fn(key: Optional[Union[str, List[str]]], aggs: list, batch_format: str,
                          data_context: DataContext,
                          _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None
                          ) -> AllToAllTransformFn:
    """planner/aggregate.py:24-98.  ``key=None``: one output block with one row (:61-63);
    else boundaries are sampled on the key and R = number of input blocks."""
    from .block import BlockExecStats
    from .exchange.aggregate_task_spec import SortAggregateTaskSpec
    from .exchange.sort_task_spec import SortTaskSpec

    assert data_context.shuffle_strategy in [ShuffleStrategy.SORT_SHUFFLE_PULL_BASED,
                                             ShuffleStrategy.SORT_SHUFFLE_PUSH_BASED,
                                             ShuffleStrategy.B200_DEVICE_EXCHANGE]
    if len(aggs) == 0:
        raise ValueError("Aggregate requires at least one aggregation")

    def fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        blocks, metadata = _flatten(refs)
        collective = (data_context.shuffle_strategy == ShuffleStrategy.B200_DEVICE_EXCHANGE
                      and _device_world() > 1)  # a rank without blocks still joins the collectives
        if len(blocks) == 0 and not collective:
            return (blocks, {})
        unified_schema = unify_block_metadata_schema(metadata)
        if unified_schema is not None or not collective:
            for agg_fn in aggs:
                agg_fn._validate(unified_schema)
        num_mappers = len(blocks)
        sort_key = SortKey(key)
        if data_context.shuffle_strategy == ShuffleStrategy.B200_DEVICE_EXCHANGE:
            from .exchange.device_exchange import DeviceExchange

            st = BlockExecStats.builder()
            out = DeviceExchange().aggregate_blocks(blocks, key, aggs, seed=int(ctx.kwargs.get("sample_seed", 0)))
            bundles, metas = _bundles_of(out, all(b.owns_blocks for b in refs), st)
            return bundles, {"map": to_stats(metadata), "reduce": to_stats(metas)}
        if key is None:
            num_outputs = 1
            boundaries = []
        else:
            num_outputs = num_mappers
            sample_bar = (ctx.sub_progress_bar_dict or {}).get(SortTaskSpec.SORT_SAMPLE_SUB_PROGRESS_BAR_NAME)
            if ctx.sub_progress_bar_dict is not None:
                assert sample_bar is not None, ctx.sub_progress_bar_dict
            boundaries = SortTaskSpec.sample_boundaries(blocks, sort_key, num_outputs, sample_bar)
        agg_spec = SortAggregateTaskSpec(boundaries=boundaries, key=sort_key, aggs=aggs, batch_format=batch_format)
        scheduler = _scheduler_for(data_context, agg_spec)
        return scheduler.execute(
            refs, num_outputs, ctx,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    return fn


def generate_randomize_blocks_fn(op: RandomizeBlocks) -> AllToAllTransformFn:
    """planner/randomize_blocks.py:12-51: the ORDER of the blocks is shuffled (seeded like
    the reference: ``random.seed(seed); random.shuffle(list)``), no row moves."""

    def fn(refs: List[RefBundle], context: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        blocks_with_metadata = []
        for ref_bundle in refs:
            blocks_with_metadata.extend(ref_bundle.blocks)
        if len(blocks_with_metadata) == 0:
            return refs, {op._name: []}
        # random.Random(seed).shuffle draws the same sequence as random.seed(seed);
        # random.shuffle on the module-level generator, without touching global state
        rng = random.Random(op._seed) if op._seed is not None else random.Random()
        input_owned = all(b.owns_blocks for b in refs)
        rng.shuffle(blocks_with_metadata)
        output, stats_list = [], []
        for block, meta in blocks_with_metadata:
            stats_list.append(meta)
            output.append(RefBundle([(block, meta)], owns_blocks=input_owned))
        return output, {op._name: stats_list}

    return fn


class SplitRepartitionTaskScheduler:
    """planner/exchange/split_repartition_task_scheduler.py:16-150: repartition WITHOUT a
    shuffle -- the rows keep their order; the concatenation of the input blocks is cut at
    the ``output_num_blocks - 1`` indices of an equal split and the pieces of each output
    are concatenated by the spec's ``reduce``."""

    def __init__(self, exchange_spec):
        self._exchange_spec = exchange_spec

    def execute(self, refs: List[RefBundle], output_num_blocks: int, ctx: TaskContext,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None
                ) -> Tuple[List[RefBundle], StatsDict]:
        from .exchange.interfaces import sub_progress_bar
        from .exchange.shuffle_task_spec import ShuffleTaskSpec

        input_num_rows = 0
        input_owned = True
        for ref_bundle in refs:
            n = ref_bundle.num_rows()
            if n is None:
                raise ValueError("Cannot split partition on blocks with unknown number of rows.")
            input_num_rows += n
            input_owned = input_owned and ref_bundle.owns_blocks
        if output_num_blocks == 1:
            indices = [input_num_rows]
        else:
            indices, cur = [], 0.0
            for _ in range(output_num_blocks - 1):
                cur += input_num_rows / output_num_blocks
                indices.append(int(cur))
        blocks = [B200ArrowBlockAccessor.for_block(b).to_arrow() for r in refs for b in r.block_refs]
        # pieces[j] = slices of the input blocks that fall into output j (_split_at_indices,
        # _internal/split.py:247): zero-copy Arrow slices
        cuts = [0] + [min(i, input_num_rows) for i in indices]
        if len(cuts) < output_num_blocks + 1:
            cuts.append(input_num_rows)
        pieces: List[List[pa.Table]] = [[] for _ in range(output_num_blocks)]
        pos = 0
        for b in blocks:
            lo, hi = pos, pos + b.num_rows
            for j in range(output_num_blocks):
                a, z = max(lo, cuts[j]), min(hi, cuts[j + 1])
                if z > a:
                    pieces[j].append(b.slice(a - lo, z - a))
            pos = hi
        reduce_bar = sub_progress_bar(ctx, ShuffleTaskSpec.SPLIT_REPARTITION_SUB_PROGRESS_BAR_NAME)
        spec = self._exchange_spec
        out_blocks, out_meta = [], []
        for j in range(output_num_blocks):
            if pieces[j]:  # only splits which contain blocks are reduced (:84-91)
                blk, meta = spec.reduce(*spec._reduce_args, *pieces[j])
                out_blocks.append(blk)
                out_meta.append(meta)
        if reduce_bar is not None:
            out_meta = reduce_bar.fetch_until_complete(out_meta)
        if len(out_blocks) < output_num_blocks:  # trailing empty blocks (:99-126)
            if not out_meta or out_meta[0].schema is None:
                raise ValueError("Cannot split partition on blocks with unknown block format.")
            for _ in range(output_num_blocks - len(out_blocks)):
                eb = empty_table()
                out_blocks.append(eb)
                out_meta.append(B200ArrowBlockAccessor.for_block(eb).get_metadata(exec_stats=None))
        output = [RefBundle([(b, m if m.size_bytes is not None else
                              B200ArrowBlockAccessor.for_block(b).get_metadata(exec_stats=m.exec_stats))],
                            owns_blocks=input_owned) for b, m in zip(out_blocks, out_meta)]
        return output, {"split": [], "reduce": out_meta}


def generate_repartition_fn(num_outputs: int, shuffle: bool, data_context: DataContext,
                            _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None
                            ) -> AllToAllTransformFn:
    """planner/repartition.py:23-83."""
    from .exchange.shuffle_task_spec import ShuffleTaskSpec

    def shuffle_repartition_fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        map_transformer = ctx.upstream_map_transformer
        upstream_map_fn = None
        if map_transformer:
            map_transformer.set_target_max_block_size(float("inf"))

            def upstream_map_fn(blocks):
                return map_transformer.apply_transform(blocks, ctx)

        shuffle_spec = ShuffleTaskSpec(ctx.target_max_block_size, random_shuffle=False,
                                       upstream_map_fn=upstream_map_fn)
        scheduler = _scheduler_for(data_context, shuffle_spec)
        return scheduler.execute(
            refs, num_outputs, ctx,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    def split_repartition_fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        shuffle_spec = ShuffleTaskSpec(ctx.target_max_block_size, random_shuffle=False)
        return SplitRepartitionTaskScheduler(shuffle_spec).execute(refs, num_outputs, ctx)

    return shuffle_repartition_fn if shuffle else split_repartition_fn


# ---- physical operators ------------------------------------------------------------------

class PhysicalOperator:
    """execution/interfaces/physical_operator.py, the part the all-to-all path uses."""

    def __init__(self, name: str, input_dependencies: List["PhysicalOperator"], data_context: DataContext,
                 target_max_block_size: Optional[int]):
        self._name = name
        self._input_dependencies = input_dependencies
        self._data_context = data_context
        self._target_max_block_size = target_max_block_size
        self._inputs_complete = not input_dependencies
        self._logical_operators: List[LogicalOperator] = []

    @property
    def name(self) -> str:
        return self._name

    @property
    def input_dependencies(self) -> List["PhysicalOperator"]:
        return self._input_dependencies

    @property
    def target_max_block_size(self) -> Optional[int]:
        return self._target_max_block_size

    @property
    def actual_target_max_block_size(self) -> int:
        """The op's own target, else the context default (physical_operator.py)."""
        if self._target_max_block_size is None:
            return self._data_context.target_max_block_size
        return self._target_max_block_size

    def set_logical_operators(self, *ops: LogicalOperator) -> None:
        self._logical_operators = list(ops)

    def all_inputs_done(self) -> None:
        self._inputs_complete = True

    def completed(self) -> bool:
        return self._inputs_complete and not self.has_next()

    def has_next(self) -> bool:
        raise NotImplementedError

    def num_outputs_total(self) -> Optional[int]:
        return None

    def num_output_rows_total(self) -> Optional[int]:
        return None


class InputDataBuffer(PhysicalOperator):
    """execution/operators/input_data_buffer.py: a source that replays given bundles."""

    def __init__(self, data_context: DataContext, input_data: List[RefBundle]):
        super().__init__("Input", [], data_context, None)
        self._input_data = list(input_data)
        self._num_output_bundles = len(self._input_data)
        self._num_output_rows = sum(b.num_rows() or 0 for b in self._input_data)

    def has_next(self) -> bool:
        return len(self._input_data) > 0

    def get_next(self) -> RefBundle:
        return self._input_data.pop(0)

    def num_outputs_total(self) -> Optional[int]:
        return self._num_output_bundles

    def num_output_rows_total(self) -> Optional[int]:
        return self._num_output_rows


class AllToAllOperator(PhysicalOperator):
    """A blocking operator that executes once its inputs are complete
    (base_physical_operator.py:42-153)."""

    def __init__(self, bulk_fn: AllToAllTransformFn, input_op: PhysicalOperator, data_context: DataContext,
                 target_max_block_size: Optional[int], num_outputs: Optional[int] = None,
                 sub_progress_bar_names: Optional[List[str]] = None, name: str = "AllToAll"):
        self._bulk_fn = bulk_fn
        self._next_task_index = 0
        self._num_outputs = num_outputs
        self._output_rows = 0
        self._sub_progress_bar_names = sub_progress_bar_names
        self._sub_progress_bar_dict: Optional[Dict[str, ProgressBar]] = None
        self._input_buffer: List[RefBundle] = []
        self._output_buffer: List[RefBundle] = []
        self._stats: StatsDict = {}
        super().__init__(name, [input_op], data_context, target_max_block_size)

    def num_outputs_total(self) -> Optional[int]:
        return self._num_outputs if self._num_outputs else self.input_dependencies[0].num_outputs_total()

    def num_output_rows_total(self) -> Optional[int]:
        return self._output_rows if self._output_rows else self.input_dependencies[0].num_output_rows_total()

    def add_input(self, refs: RefBundle, input_index: int = 0) -> None:
        assert not self.completed()
        assert input_index == 0, input_index
        self._input_buffer.append(refs)

    def all_inputs_done(self) -> None:
        ctx = TaskContext(task_idx=self._next_task_index, sub_progress_bar_dict=self._sub_progress_bar_dict,
                          target_max_block_size=self.actual_target_max_block_size)
        self._output_buffer, self._stats = self._bulk_fn(self._input_buffer, ctx)
        self._next_task_index += 1
        self._input_buffer.clear()
        super().all_inputs_done()

    def has_next(self) -> bool:
        return len(self._output_buffer) > 0

    def get_next(self) -> RefBundle:
        bundle = self._output_buffer.pop(0)
        self._output_rows += bundle.num_rows()
        return bundle

    def get_stats(self) -> StatsDict:
        return self._stats

    def get_transformation_fn(self) -> AllToAllTransformFn:
        return self._bulk_fn

    def progress_str(self) -> str:
        return f"{self.num_output_rows_total() or 0} rows output"

    def initialize_sub_progress_bars(self, position: int) -> int:
        if self._sub_progress_bar_names is not None:
            self._sub_progress_bar_dict = {}
            for name in self._sub_progress_bar_names:
                bar = ProgressBar(name, self.num_output_rows_total() or 1, unit="row", position=position)
                bar.set_description(f"  *- {name}")
                self._sub_progress_bar_dict[name] = bar
                position += 1
            return len(self._sub_progress_bar_dict)
        return 0

    def close_sub_progress_bars(self) -> None:
        if self._sub_progress_bar_dict is not None:
            for sub_bar in self._sub_progress_bar_dict.values():
                sub_bar.close()

    def supports_fusion(self) -> bool:
        return True


def plan_all_to_all_op(op: AbstractAllToAll, physical_children: List[PhysicalOperator],
                       data_context: DataContext) -> AllToAllOperator:
    """planner/plan_all_to_all_op.py:23-105: shuffle-type operators get the 1 GiB
    ``target_shuffle_max_block_size``; RandomizeBlocks and split-repartition inherit."""
    assert len(physical_children) == 1
    input_physical_dag = physical_children[0]
    target_max_block_size = None
    debug_limit = data_context.get_config("debug_limit_shuffle_execution_to_num_blocks", None)
    if isinstance(op, RandomizeBlocks):
        fn = generate_randomize_blocks_fn(op)
    elif isinstance(op, RandomShuffle):
        fn = generate_random_shuffle_fn(data_context, op._seed, op._num_outputs, op._ray_remote_args, debug_limit)
        target_max_block_size = data_context.target_shuffle_max_block_size
    elif isinstance(op, Repartition):
        if op._shuffle:
            target_max_block_size = data_context.target_shuffle_max_block_size
        fn = generate_repartition_fn(op._num_outputs, op._shuffle, data_context,
                                     debug_limit if op._shuffle else None)
    elif isinstance(op, Sort):
        fn = generate_sort_fn(op._sort_key, op._batch_format, data_context, debug_limit)
        target_max_block_size = data_context.target_shuffle_max_block_size
    elif isinstance(op, Aggregate):
        fn = generate_aggregate_fn(op._key, op._aggs, op._batch_format, data_context, debug_limit)
        target_max_block_size = data_context.target_shuffle_max_block_size
    else:
        raise ValueError(f"Found unknown logical operator during planning: {op}")
    phys = AllToAllOperator(fn, input_physical_dag, data_context, target_max_block_size=target_max_block_size,
                            num_outputs=op._num_outputs, sub_progress_bar_names=op._sub_progress_bar_names,
                            name=op.name)
    phys.set_logical_operators(op)
    return phys


def execute_all_to_all(op: AbstractAllToAll, input_bundles: List[RefBundle],
                       data_context: Optional[DataContext] = None) -> Tuple[List[RefBundle], StatsDict]:
    """Plan ``op`` over in-memory bundles and run it the way the streaming executor drives
    an AllToAllOperator: feed every input bundle, signal completion, drain the outputs."""
    data_context = data_context or DataContext.get_current()
    src = InputDataBuffer(data_context, input_bundles)
    phys = plan_all_to_all_op(op, [src], data_context)
    phys.initialize_sub_progress_bars(0)
    while src.has_next():
        phys.add_input(src.get_next(), 0)
    phys.all_inputs_done()
    out = []
    while phys.has_next():
        out.append(phys.get_next())
    phys.close_sub_progress_bars()
    return out, phys.get_stats()

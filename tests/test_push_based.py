"""Push-based (Exoshuffle) schedule: the schedule arithmetic against the reference's own
invariants (python/ray/data/tests/test_sort.py:356-470) and -- in the build container --
against the reference's functions themselves; and map/merge/reduce rounds on the engine
giving the same blocks as the pull-based schedule (sort, aggregate) / the same multiset
(shuffle)."""
import os
from collections import defaultdict

import numpy as np
import pyarrow as pa
import pytest

import _golden as G
from cluster_anywhere_b200 import Count, Mean, SortKey, Std, Sum
from cluster_anywhere_b200.exchange import (PullBasedShuffleTaskScheduler, PushBasedShuffleTaskScheduler,
                                            ShuffleTaskSpec, SortAggregateTaskSpec, SortTaskSpec)
from cluster_anywhere_b200.exchange.push_based import MergeSchedule, compute_push_schedule

_GRID = ([(20, 3, {"node1": c}) for c in range(1, 20)] +
         [(20, 3, {"node1": 100}), (100, 3, {"node1": 10, "node2": 10, "node3": 10}),
          (100, 10, {"node1": 10, "node2": 10, "node3": 10}),
          (1000, 2, {f"node{i}": 16 for i in range(20)}), (260, 2, {"node1": 128}), (1, 2, {"node1": 128})] +
         [(n, f, cfg) for cfg in ({"node1": 10}, {"node1": 10, "node2": 10})
          for n, f in ((100, 1), (100, 1.3), (100, 1.6), (100, 1.75), (100, 2), (1, 1.2), (2, 1.2))])


@pytest.mark.parametrize("num_input_blocks,merge_factor,cpus", _GRID)
def test_schedule_invariants(num_input_blocks, merge_factor, cpus):
    num_cpus = sum(cpus.values())
    st = compute_push_schedule(cpus, num_input_blocks, merge_factor, num_input_blocks)
    ms = st.merge_schedule
    assert st.num_rounds * st.num_map_tasks_per_round >= num_input_blocks
    assert st.num_map_tasks_per_round + ms.num_merge_tasks_per_round <= max(num_cpus, 2)
    if st.num_map_tasks_per_round > merge_factor:
        actual = st.num_map_tasks_per_round / ms.num_merge_tasks_per_round
        assert actual - 1 <= merge_factor <= actual + 1
    else:
        assert ms.num_merge_tasks_per_round == 1
    per_node = defaultdict(int)
    for node in st.merge_task_placement[:ms.num_merge_tasks_per_round]:
        per_node[node] += 1
    assert max(per_node.values()) <= min(per_node.values()) + 1
    counts = [ms.get_num_reducers_per_merge_idx(i) for i in range(ms.num_merge_tasks_per_round)]
    assert all(c in (max(counts), max(counts) - 1) and c > 0 for c in counts)
    seen = defaultdict(int)
    order = list(ms.round_robin_reduce_idx_iterator())
    assert sorted(order) == list(range(ms.output_num_blocks))
    for j in order:
        seen[ms.get_merge_idx_for_reducer_idx(j)] += 1
    assert [seen[i] for i in range(ms.num_merge_tasks_per_round)] == counts


@pytest.mark.skipif(not os.path.isdir("/root/reference/python/ray"), reason="reference tree not present")
def test_schedule_equals_the_reference():
    from oracle import ref_shim

    ref_shim.load()
    from ray.data._internal.planner.exchange.push_based_shuffle_task_scheduler import (
        PushBasedShuffleTaskScheduler as RefSched, _MergeTaskSchedule)

    for n, f, cpus in _GRID:
        for n_out in (n, max(1, n // 3), n * 2 + 1):
            a = compute_push_schedule(cpus, n, f, n_out)
            b = RefSched._compute_shuffle_schedule(cpus, n, f, n_out)
            assert (a.num_rounds, a.num_map_tasks_per_round) == (b.num_rounds, b.num_map_tasks_per_round)
            assert a.merge_schedule.num_merge_tasks_per_round == b.merge_schedule.num_merge_tasks_per_round
            assert a.get_estimated_num_refs() == b.get_estimated_num_refs()
            m = b.merge_schedule.num_merge_tasks_per_round
            assert [a.merge_schedule.get_num_reducers_per_merge_idx(i) for i in range(m)] == \
                [b.merge_schedule.get_num_reducers_per_merge_idx(i) for i in range(m)]
            assert list(a.merge_schedule.round_robin_reduce_idx_iterator()) == \
                list(b.merge_schedule.round_robin_reduce_idx_iterator())
    for r, m in ((10, 3), (3, 10), (7, 7), (1, 4), (100, 9)):
        a, b = MergeSchedule(r, m), _MergeTaskSchedule(r, m)
        assert a.num_merge_tasks_per_round == b.num_merge_tasks_per_round
        assert [a.get_merge_idx_for_reducer_idx(j) for j in range(r)] == \
            [b.get_merge_idx_for_reducer_idx(j) for j in range(r)]


@pytest.mark.parametrize("case", G.cases("sort"), ids=lambda c: c["name"])
def test_push_sort_equals_pull_and_golden(engine, case):
    blocks = [G.load(f) for f in case["inputs"]]
    exp = [G.load(f) for f in case["outputs"]]
    sk = SortKey(case["columns"], case["descending"])
    spec = SortTaskSpec(G.boundaries(case), sk)
    for cpus in ({"n": 3}, {"n": 16}, {"a": 2, "b": 2}):
        out, stats = PushBasedShuffleTaskScheduler(spec).execute(blocks, len(exp), num_cpus_per_node_map=cpus)
        assert set(stats) == {"map", "merge", "reduce"}
        assert len(out) == len(exp)
        for o, e in zip(out, exp):
            assert o.num_rows == e.num_rows
            if e.num_rows:
                assert G.tables_equal(o, e)


def test_push_aggregate_and_shuffle(engine):
    case = G.cases("aggregate")[0]
    blocks = [G.load(f) for f in case["inputs"]]
    aggs = [Count(), Sum("v"), Mean("v"), Std("v"), Sum("w")]
    spec = SortAggregateTaskSpec(G.boundaries(case), SortKey(case["key"]), aggs)
    pull, _ = PullBasedShuffleTaskScheduler(spec).execute(blocks, len(case["outputs"]))
    push, stats = PushBasedShuffleTaskScheduler(spec).execute(blocks, len(case["outputs"]),
                                                             num_cpus_per_node_map={"n": 3})
    assert set(stats) == {"map", "merge", "reduce"}
    from test_aggregate import assert_tables_close

    for a, b in zip(push, pull):
        assert_tables_close(a, b, exact=(case["key"],))
    sc = G.cases("shuffle")[0]
    sblocks = [G.load(f) for f in sc["inputs"]]
    sspec = ShuffleTaskSpec(1 << 30, True, sc["seed"])
    out, _ = PushBasedShuffleTaskScheduler(sspec).execute(sblocks, sc["num_outputs"], num_cpus_per_node_map={"n": 3})
    got = sorted(map(repr, pa.concat_tables(out).to_pylist()))
    assert got == sorted(map(repr, pa.concat_tables(sblocks).to_pylist()))
    out2, _ = PushBasedShuffleTaskScheduler(sspec).execute(sblocks, sc["num_outputs"], num_cpus_per_node_map={"n": 3})
    assert all(G.tables_equal(a, b) for a, b in zip(out, out2))

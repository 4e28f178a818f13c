"""The task-group optimizer mirror against the reference's own tests
(/root/reference/ballista/core/src/physical_optimizer/task_group.rs:335-446): a CoalesceTasksExec goes on top of a leaf,
a Partial aggregate is pushed below it, and the stage's writer takes the group's partitions - which yields exactly the two
stage shapes `GpuExecutionEngine` takes to the device. No GPU."""
import os
import sys

import pyarrow as pa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Scan:
    def __init__(self, schema, n):
        self._schema, self.n = schema, n

    def schema(self):
        return self._schema

    def children(self):
        return []

    def execute(self, p):
        return iter(())


def q1(bb):
    sys.path.insert(0, ROOT)
    import bench

    gi, aggs, filt = bench.q1_plan()
    return gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs], filt, bench.lineitem_schema()


def test_insert_coalesce_on_top_of_the_leaf(bb):
    # test_optimizer_insert_coalesce_on_top_of_the_leaf (:335-355)
    reader = bb.GpuShuffleReaderExec([], pa.schema([]), None)
    out = bb.OptimizeTaskGroup([])._transform_down(reader)
    assert isinstance(out, bb.CoalesceTasksExec) and out.children() == [reader] and reader.children() == []


def test_push_partial_agg(bb, tmp_path):
    # test_optimizer_push_partial_agg (:408-446): writer -> CoalesceTasksExec -> AggregateExec(Partial) -> scan
    gi, aggs, filt, schema = q1(bb)
    scan = Scan(schema, 10)
    agg = bb.GpuAggregateExecNode(bb.AggregateMode.Partial, gi, aggs, schema, filt, inputs=[scan])
    stage = bb.GpuShuffleWriterExec.try_new("job", 1, [], agg, "", bb.Partitioning.Hash([0, 1], 16), None)
    optimized = bb.OptimizeTaskGroup(range(10)).optimize(stage)
    assert isinstance(optimized, bb.GpuShuffleWriterExec) and optimized.partitions == list(range(10))
    (child,) = optimized.children()
    assert isinstance(child, bb.CoalesceTasksExec) and child.partitions == list(range(10))
    (child,) = child.children()
    assert isinstance(child, bb.GpuAggregateExecNode) and child.mode == bb.AggregateMode.Partial
    assert child.children() == [scan]
    # ... which is the map stage the engine takes to the device
    engine = bb.GpuExecutionEngine(ctx=object())
    exec_ = engine.create_query_stage_exec("job", 1, optimized, str(tmp_path))
    assert isinstance(exec_, bb.GpuQueryStageExec) and exec_.partitions == list(range(10)) and exec_.source is scan


def test_reduce_stage_keeps_the_fan_in_above_the_final_aggregate(bb, tmp_path):
    # task_group.rs:143-151: the coalesce goes above a FinalPartitioned aggregate and the walk stops there, so the
    # CoalesceBatchesExec + reader below it stay per input partition (planner.rs:405-413 plans them in that order)
    gi, aggs, _, schema = q1(bb)
    reader = bb.GpuShuffleReaderExec([[], [], [], []], None, None)
    batches = bb.CoalesceBatchesExec(reader, None)
    agg = bb.GpuAggregateExecNode(bb.AggregateMode.FinalPartitioned, gi, aggs, schema, None, inputs=[batches])
    stage = bb.GpuShuffleWriterExec.try_new("job", 2, [], agg, "", None, None)
    optimized = bb.OptimizeTaskGroup([1, 3]).optimize(stage)
    assert optimized.partitions == [1, 3]
    (fan_in,) = optimized.children()
    assert isinstance(fan_in, bb.CoalesceTasksExec) and fan_in.partitions == [1, 3]
    assert fan_in.children() == [agg] and agg.children() == [batches] and batches.children() == [reader]
    exec_ = bb.GpuExecutionEngine(ctx=object()).create_query_stage_exec("job", 2, optimized, str(tmp_path))
    assert isinstance(exec_, bb.GpuQueryStageExec) and exec_.partitions == [1, 3] and exec_.source is batches


def test_mapping_chain_is_pushed_below_the_fan_in(bb):
    # CoalesceBatchesExec and a Partial aggregate above a leaf both end up below ONE CoalesceTasksExec (:77-110)
    gi, aggs, filt, schema = q1(bb)
    scan = Scan(schema, 4)
    agg = bb.GpuAggregateExecNode(bb.AggregateMode.Partial, gi, aggs, schema, filt, inputs=[bb.CoalesceBatchesExec(scan, None)])
    stage = bb.GpuShuffleWriterExec.try_new("job", 1, [], agg, "", bb.Partitioning.Hash([0, 1], 16), None)
    optimized = bb.OptimizeTaskGroup([0, 1]).optimize(stage)
    (fan_in,) = optimized.children()
    assert isinstance(fan_in, bb.CoalesceTasksExec)
    (a,) = fan_in.children()
    assert isinstance(a, bb.GpuAggregateExecNode)
    (b,) = a.children()
    assert isinstance(b, bb.CoalesceBatchesExec) and b.children() == [scan]

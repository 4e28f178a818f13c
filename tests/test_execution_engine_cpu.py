"""Host logic of the stage-execution hook: which decoded stage plans `GpuExecutionEngine` takes to the
device and which go to the fallback engine, mirroring /root/reference/ballista/executor/src/execution_engine.rs:63-97
(the writer is re-created with the executor's work_dir; a plan whose root is no shuffle writer is an error). No GPU."""
import os
import sys

import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Scan:
    """A leaf with three input partitions."""

    def __init__(self, schema):
        self._schema = schema

    def schema(self):
        return self._schema

    def execute(self, partition):
        return iter(())


def q1_stage(bb, mode=None, coalesce=True, partitioning="hash", fns=None):
    sys.path.insert(0, ROOT)
    import bench

    gi, aggs, filt = bench.q1_plan()
    if fns:
        aggs = [(fn, a, n) for fn, (_, a, n) in zip(fns, aggs)]
    schema = bench.lineitem_schema()
    node = bb.GpuAggregateExecNode(bb.AggregateMode.Partial if mode is None else mode, gi,
                                   [(bb.AggregateFunction.parse(f), bb.expr.col(a) if isinstance(a, int) else a, n) for f, a, n in aggs],
                                   schema, filt, inputs=[Scan(schema)])
    child = bb.CoalesceTasksExec(node, [4, 5, 6]) if coalesce else node
    part = {"hash": bb.Partitioning.Hash([0, 1], 16), "none": None}[partitioning]
    return bb.GpuShuffleWriterExec.try_new("job", 1, [4, 5, 6], child, "", part, None), node


def test_hot_path_map_stage_goes_to_the_device(bb, tmp_path):
    plan, node = q1_stage(bb)
    engine = bb.GpuExecutionEngine()
    stage = engine.create_query_stage_exec("job7", 2, plan, str(tmp_path))
    assert isinstance(stage, bb.GpuQueryStageExec) and engine.last_decision == "gpu"
    # the writer is re-created with the executor's work_dir and the ids of the call (execution_engine.rs:77-87)
    w = stage.shuffle_writer
    assert (w.job_id, w.stage_id, w.work_dir, w.partitions) == ("job7", 2, str(tmp_path), [4, 5, 6])
    assert w.shuffle_output_partitioning.partition_count() == 16 and w.shuffle_output_partitioning.exprs == [0, 1]
    assert stage.aggregate is node and stage.partitions == [4, 5, 6]
    assert "GpuAggregateExec(Partial" in str(stage) and "Hash([0, 1], 16)" in str(stage)
    # without the CoalesceTasksExec the partitions come with execute_query_stage
    plan, _ = q1_stage(bb, coalesce=False)
    stage = engine.create_query_stage_exec("job7", 2, plan, str(tmp_path))
    assert isinstance(stage, bb.GpuQueryStageExec) and stage.partitions is None


@pytest.mark.parametrize("kw, why", [
    (dict(partitioning="none"), "does not hash-repartition"),
    (dict(mode=1), "neither Partial nor FinalPartitioned"),
    (dict(fns=["min"]), "other than SUM / AVG / COUNT"),
])
def test_other_stages_go_to_the_fallback_engine(bb, tmp_path, kw, why):
    plan, _ = q1_stage(bb, **kw)
    engine = bb.GpuExecutionEngine()
    stage = engine.create_query_stage_exec("j", 1, plan, str(tmp_path))
    assert isinstance(stage, bb.DefaultQueryStageExec) and why in engine.last_decision
    assert stage.shuffle_writer.work_dir == str(tmp_path) and stage.shuffle_writer.plan is plan.plan

    class Theirs:
        def create_query_stage_exec(self, *a):
            return ("theirs",) + a
    got = bb.GpuExecutionEngine(fallback=Theirs()).create_query_stage_exec("j", 1, plan, str(tmp_path))
    assert got == ("theirs", "j", 1, plan, str(tmp_path), None)


def test_child_that_is_a_plain_stream_and_partitioning_over_state_columns_fall_back(bb, tmp_path):
    schema = pa.schema([("a", pa.int64())])
    plan = bb.GpuShuffleWriterExec.try_new("j", 1, [0], [], "", bb.Partitioning.Hash([0], 4), None, input_schema=schema)
    engine = bb.GpuExecutionEngine()
    assert isinstance(engine.create_query_stage_exec("j", 1, plan, str(tmp_path)), bb.DefaultQueryStageExec)
    assert "not an aggregate" in engine.last_decision
    plan, _ = q1_stage(bb)
    plan.shuffle_output_partitioning = bb.Partitioning.Hash([0, 2], 16)  # column 2 of the state batch is a sum, not a key
    assert isinstance(engine.create_query_stage_exec("j", 1, plan, str(tmp_path)), bb.DefaultQueryStageExec)
    assert "group columns" in engine.last_decision


def test_root_that_is_no_shuffle_writer_is_an_error(bb, tmp_path):
    _, node = q1_stage(bb)
    for engine in (bb.GpuExecutionEngine(), bb.DefaultExecutionEngine()):
        with pytest.raises(bb.BgpuError, match="Plan passed to new_query_stage_exec is not a ShuffleWriterExec"):
            engine.create_query_stage_exec("j", 1, node, str(tmp_path))


def test_stage_plan_survives_the_codec_and_still_matches(bb, tmp_path):
    """scheduler -> executor: the optimized stage plan is encoded node by node (children ride as
    PhysicalExtensionNode.inputs), decoded with work_dir "" and handed to the engine (executor_server.rs:383-397)."""
    plan, node = q1_stage(bb, coalesce=False)
    codec = bb.GpuPhysicalCodec()
    scan = node.inputs[0]
    agg = codec.try_decode(codec.try_encode(node), [scan])
    writer = codec.try_decode(codec.try_encode(plan), [agg])
    assert writer.work_dir == "" and writer.plan is agg and agg.inputs == [scan]
    stage = bb.GpuExecutionEngine().create_query_stage_exec("job", 1, writer, str(tmp_path))
    assert isinstance(stage, bb.GpuQueryStageExec) and stage.source is scan
    assert stage.aggregate.group_by == node.group_by and stage.shuffle_writer.work_dir == str(tmp_path)


def test_stage_with_no_input_partitions_writes_nothing(bb, tmp_path):
    plan, _ = q1_stage(bb, coalesce=False)
    stage = bb.GpuExecutionEngine(ctx=object()).create_query_stage_exec("job", 1, plan, str(tmp_path))
    assert stage.execute_query_stage([]) == [] and not os.listdir(tmp_path)
    assert stage.collect_plan_metrics()[0]["aggregate_input_rows"] == 0


def test_reduce_stage_goes_to_the_device(bb, tmp_path):
    """ShuffleWriterExec(None) -> CoalesceTasksExec -> AggregateExec(FinalPartitioned) -> CoalesceBatchesExec -> reader
    (planner.rs:405-413 + task_group.rs:143-151: the coalesce sits above a Final aggregate)."""
    sys.path.insert(0, ROOT)
    import bench

    gi, aggs, _ = bench.q1_plan()
    schema = bench.lineitem_schema()
    reader = bb.GpuShuffleReaderExec([[], [], []], None, None)
    source = bb.CoalesceBatchesExec(reader, None)
    assert source.children() == [reader]
    node = bb.GpuAggregateExecNode(bb.AggregateMode.FinalPartitioned, gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs],
                                   schema, None, inputs=[source])
    plan = bb.GpuShuffleWriterExec.try_new("job", 2, [0, 1, 2], bb.CoalesceTasksExec(node, [0, 1, 2]), "", None, None)
    engine = bb.GpuExecutionEngine(ctx=object())
    stage = engine.create_query_stage_exec("job", 2, plan, str(tmp_path))
    assert isinstance(stage, bb.GpuQueryStageExec) and engine.last_decision == "gpu"
    assert stage.source is source and stage.partitions == [0, 1, 2] and "GpuShuffleWriterExec(None)" in str(stage)
    assert "FinalPartitioned" in str(stage)


def test_reduce_stage_plumbing_with_a_stand_in_aggregate(bb, tmp_path, monkeypatch):
    """execute_query_stage's host plumbing without a device: the aggregate is replaced by a stand-in that hands back
    host batches, so that what is checked is the walk over the task group's partitions, the skipping of partitions
    nobody wrote to, the writer re-creation with the stage's schema, the un-partitioned `data.arrow`
    (shuffle_writer.rs:258-326) and the metrics."""
    schema = pa.schema([("k", pa.int64()), ("v", pa.int64())])
    parts = {0: [pa.record_batch({"k": [1, 2], "v": [10, 20]}, schema=schema)], 1: [],
             2: [pa.record_batch({"k": [3], "v": [30]}, schema=schema), pa.record_batch({"k": [4], "v": [40]}, schema=schema)]}

    class Source:
        def execute(self, p):
            return iter(parts[p])

    class StandIn:
        hot_kernel_ms, last_path = 0.0, "stand-in"

        def __init__(self):
            self.rows = []

        def update(self, b):
            self.rows.append(b)

        def emit(self):
            return pa.Table.from_batches(self.rows).combine_chunks().to_batches()[0]

        def close(self):
            pass
    node = bb.GpuAggregateExecNode(bb.AggregateMode.FinalPartitioned, [0], [(bb.AggregateFunction.Sum, bb.expr.col(1), "v")], schema,
                                   inputs=[Source()])
    monkeypatch.setattr(bb.GpuAggregateExecNode, "instantiate", lambda self, ctx: StandIn())
    for merge, n_batches in ((True, 1), (False, 2)):
        plan = bb.GpuShuffleWriterExec.try_new("job", 2, [0, 1, 2], bb.CoalesceTasksExec(node, [0, 1, 2]), "", None, None)
        stage = bb.GpuExecutionEngine(ctx=object(), merge_partitions=merge).create_query_stage_exec("job", 2, plan, str(tmp_path))
        (w,) = stage.execute_query_stage([0, 1, 2])
        assert (w.partitions, w.output_partition, w.num_rows, w.num_batches) == ([0, 1, 2], 0, 4, n_batches)
        assert w.path.startswith(os.path.join(str(tmp_path), "job", "2")) and w.path.endswith("data.arrow")
        with pa.ipc.open_stream(w.path) as r:
            got = r.read_all()
        assert got.sort_by("k").to_pydict() == {"k": [1, 2, 3, 4], "v": [10, 20, 30, 40]}
        assert stage.schema().equals(schema)
        m_agg, m_write = stage.collect_plan_metrics()
        assert m_agg["aggregate_input_rows"] == 4 == m_agg["aggregate_output_rows"] and m_agg["last_path"] == "stand-in"
        assert m_write["input_rows"] == 4 == m_write["output_rows"]

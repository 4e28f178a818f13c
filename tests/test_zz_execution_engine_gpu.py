"""The stage-execution hook end to end on the device: two map tasks of TPC-H Q1 as the scheduler ships
them - ShuffleWriterExec(Hash(keys, 4)) -> CoalesceTasksExec(3 partitions) -> AggregateExec(Partial) -> scan, encoded and
decoded by GpuPhysicalCodec, handed to GpuExecutionEngine.create_query_stage_exec with the executor's work_dir
(/root/reference/ballista/executor/src/executor_server.rs:383-397) - then the reduce stage over the files they wrote.
The union of the reduce outputs must equal the oracle's Partial -> Final, bit-exact, whether the task group's partitions
feed one Partial aggregate or one each."""
import os
import sys

import numpy as np
import pyarrow as pa
import pytest

from helpers import expr_to_tree, lineitem_batch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Scan:
    def __init__(self, schema, partitions):
        self._schema, self.partitions = schema, partitions

    def schema(self):
        return self._schema

    def execute(self, partition):
        return iter(self.partitions[partition])


TASKS = ([0, 1, 2], [3, 4, 5])
N_OUT = 4


def q1_inputs():
    sys.path.insert(0, ROOT)
    import bench

    rng = np.random.default_rng(23)
    # 6 input partitions of the map stage, 3 per task
    data = {p: [lineitem_batch(rng, int(n)) for n in rng.integers(1, 3000, 2)] for p in range(6)}
    return bench.lineitem_schema(), bench.q1_plan(), data


def run_map_stages(bb, ctx, work_dir, schema, plan3, data, merge_partitions):
    """Both map tasks through codec + engine; returns (PartitionLocations per reduce partition, state schema)."""
    gi, aggs, filt = plan3
    scan = Scan(schema, data)
    codec = bb.GpuPhysicalCodec()
    engine = bb.GpuExecutionEngine(ctx, merge_partitions=merge_partitions)
    locations = [[] for _ in range(N_OUT)]
    state_schema = None
    for task_partitions in TASKS:
        task_partitions = list(task_partitions)
        node = bb.GpuAggregateExecNode(bb.AggregateMode.Partial, gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs],
                                       schema, filt, inputs=[scan])
        plan = bb.GpuShuffleWriterExec.try_new("q1", 1, task_partitions, bb.CoalesceTasksExec(node, task_partitions), "",
                                               bb.Partitioning.Hash(list(range(len(gi))), N_OUT), None)
        # what crosses the wire: the two GPU nodes as bytes, decoded on the executor with their children next to them
        agg = codec.try_decode(codec.try_encode(node), [scan])
        writer = codec.try_decode(codec.try_encode(plan), [bb.CoalesceTasksExec(agg, task_partitions)])
        assert writer.work_dir == ""
        stage = engine.create_query_stage_exec("q1", 1, writer, work_dir)
        assert isinstance(stage, bb.GpuQueryStageExec), engine.last_decision
        written = stage.execute_query_stage(task_partitions)
        state_schema = stage.schema()
        m_agg, m_write = stage.collect_plan_metrics()
        assert m_agg["aggregate_input_rows"] == sum(b.num_rows for p in task_partitions for b in data[p])
        assert m_write["input_rows"] == m_agg["aggregate_output_rows"] == m_write["output_rows"]
        assert m_agg["last_path"] != ""
        # the GPU node's own time metrics: host batches were uploaded, kernels ran, partition files were copied out
        assert m_write["h2d_time"] > 0 and m_write["kernel_time"] > 0 and m_write["d2h_time"] > 0, m_write
        assert m_write["h2d_time"] + m_write["kernel_time"] + m_write["d2h_time"] < 60
        for w in written:
            assert w.partitions == task_partitions and w.path.startswith(work_dir) and os.path.getsize(w.path) > w.num_bytes
            locations[w.output_partition].append(bb.PartitionLocation.from_shuffle_write("q1", 1, w))
    assert [f.name for f in state_schema][: len(gi)] == [schema.field(i).name for i in gi]
    return locations, state_schema


def oracle_answer(oracle, plan3, data):
    gi, aggs, filt = plan3
    o_aggs = [(f, expr_to_tree(a), nm) for f, a, nm in aggs]
    states = []
    for task_partitions in TASKS:
        op = oracle.GroupedAggregate("partial", gi, o_aggs, filter=expr_to_tree(filt))
        for p in task_partitions:
            for b in data[p]:
                op.update(b)
        states.append(op.emit())
    arg_types = [None if f == "count" else states[0].schema.field(f"{nm}[sum]").type for f, _, nm in aggs]
    arg_types = [None if t is None else pa.decimal128(t.precision - 10, t.scale) for t in arg_types]
    of = oracle.GroupedAggregate("final", list(range(len(gi))), [(f, 0, nm) for f, _, nm in aggs], arg_types=arg_types)
    for s in states:
        of.update(s)
    return pa.Table.from_batches([of.emit()])


def same_groups(got, want, n_keys):
    keys = [(n_, "ascending") for n_ in got.schema.names[:n_keys]]
    assert got.sort_by(keys).equals(want.sort_by(keys))


@pytest.mark.parametrize("merge_partitions", [True, False])
def test_q1_map_stages_through_the_engine(bb, ctx, oracle, tmp_path, merge_partitions):
    schema, plan3, data = q1_inputs()
    gi, aggs, _ = plan3
    locations, state_schema = run_map_stages(bb, ctx, str(tmp_path), schema, plan3, data, merge_partitions)
    reader = bb.GpuShuffleReaderExec(locations, state_schema, ctx)
    finals = []
    for p in range(N_OUT):
        fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, aggs, schema)
        seen = 0
        for b in bb.CoalesceBatchesExec(reader.execute(p), ctx):
            fin.update(b)
            seen += b.num_rows
        if seen:
            finals.append(fin.emit().download())
    same_groups(pa.Table.from_batches(finals), oracle_answer(oracle, plan3, data), len(gi))


def test_whole_q1_job_through_the_engine(bb, ctx, oracle, tmp_path):
    """Map stages as above, then the reduce stage as one task group over all four reduce partitions:
    ShuffleWriterExec(None) -> CoalesceTasksExec -> AggregateExec(FinalPartitioned) -> CoalesceBatchesExec ->
    GpuShuffleReaderExec; the job's result is the `data.arrow` the final stage writes (shuffle_writer.rs:258-326), read
    back here with Arrow's own reader."""
    schema, plan3, data = q1_inputs()
    gi, aggs, _ = plan3
    locations, state_schema = run_map_stages(bb, ctx, str(tmp_path), schema, plan3, data, True)
    reduce_partitions = list(range(N_OUT))
    source = bb.CoalesceBatchesExec(bb.GpuShuffleReaderExec(locations, state_schema, ctx), ctx)
    node = bb.GpuAggregateExecNode(bb.AggregateMode.FinalPartitioned, gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs],
                                   schema, None, inputs=[source])
    plan = bb.GpuShuffleWriterExec.try_new("q1", 2, reduce_partitions, bb.CoalesceTasksExec(node, reduce_partitions), "", None, None)
    engine = bb.GpuExecutionEngine(ctx)
    stage = engine.create_query_stage_exec("q1", 2, plan, str(tmp_path))
    assert isinstance(stage, bb.GpuQueryStageExec), engine.last_decision
    written = stage.execute_query_stage(reduce_partitions)
    assert len(written) == 1 and written[0].path.endswith("data.arrow") and written[0].output_partition == 0
    assert written[0].partitions == reduce_partitions
    with pa.ipc.open_stream(written[0].path) as r:
        got = r.read_all()
    assert written[0].num_rows == got.num_rows and written[0].num_batches == 1
    same_groups(got, oracle_answer(oracle, plan3, data), len(gi))

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


@pytest.mark.parametrize("merge_partitions", [True, False])
def test_q1_map_stages_through_the_engine(bb, ctx, oracle, tmp_path, merge_partitions):
    sys.path.insert(0, ROOT)
    import bench

    rng = np.random.default_rng(23)
    schema = bench.lineitem_schema()
    gi, aggs, filt = bench.q1_plan()
    n_out = 4
    # 6 input partitions of the stage, 3 per task
    data = {p: [lineitem_batch(rng, int(n)) for n in rng.integers(1, 3000, 2)] for p in range(6)}
    scan = Scan(schema, data)
    codec = bb.GpuPhysicalCodec()
    engine = bb.GpuExecutionEngine(ctx, merge_partitions=merge_partitions)
    locations = [[] for _ in range(n_out)]
    state_schema = None
    for task_partitions in ([0, 1, 2], [3, 4, 5]):
        node = bb.GpuAggregateExecNode(bb.AggregateMode.Partial, gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs],
                                       schema, filt, inputs=[scan])
        plan = bb.GpuShuffleWriterExec.try_new("q1", 1, task_partitions, bb.CoalesceTasksExec(node, task_partitions), "",
                                               bb.Partitioning.Hash(list(range(len(gi))), n_out), None)
        # what crosses the wire: the two GPU nodes as bytes, decoded on the executor with their children next to them
        agg = codec.try_decode(codec.try_encode(node), [scan])
        writer = codec.try_decode(codec.try_encode(plan), [bb.CoalesceTasksExec(agg, task_partitions)])
        assert writer.work_dir == ""
        stage = engine.create_query_stage_exec("q1", 1, writer, str(tmp_path))
        assert isinstance(stage, bb.GpuQueryStageExec), engine.last_decision
        written = stage.execute_query_stage(task_partitions)
        state_schema = stage.schema()
        m_agg, m_write = stage.collect_plan_metrics()
        assert m_agg["aggregate_input_rows"] == sum(b.num_rows for p in task_partitions for b in data[p])
        assert m_write["input_rows"] == m_agg["aggregate_output_rows"] == m_write["output_rows"]
        assert m_agg["last_path"] != ""
        for w in written:
            assert w.partitions == task_partitions and w.path.startswith(str(tmp_path)) and os.path.getsize(w.path) > w.num_bytes
            locations[w.output_partition].append(bb.PartitionLocation.from_shuffle_write("q1", 1, w))
    assert [f.name for f in state_schema][: len(gi)] == [schema.field(i).name for i in gi]

    reader = bb.GpuShuffleReaderExec(locations, state_schema, ctx)
    finals = []
    for p in range(n_out):
        fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, aggs, schema)
        seen = 0
        for b in bb.CoalesceBatchesExec(reader.execute(p), ctx):
            fin.update(b)
            seen += b.num_rows
        if seen:
            finals.append(fin.emit().download())
    got = pa.Table.from_batches(finals)

    o_aggs = [(f, expr_to_tree(a), nm) for f, a, nm in aggs]
    states = []
    for task_partitions in ([0, 1, 2], [3, 4, 5]):
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
    want = pa.Table.from_batches([of.emit()])
    keys = [(n_, "ascending") for n_ in got.schema.names[: len(gi)]]
    assert got.sort_by(keys).equals(want.sort_by(keys))

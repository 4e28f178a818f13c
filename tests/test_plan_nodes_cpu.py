"""The ExecutionPlan answers of the operator mirrors (SURVEY.md §8b "Plan-node trait"), against what the reference's nodes
answer: ShuffleWriterExec (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:459-516, 205-223),
ShuffleReaderExec (shuffle_reader.rs:203-224), CoalesceTasksExec (coalesce_tasks.rs:105-128). No GPU."""
import pyarrow as pa
import pytest


class Leaf:
    def __init__(self, schema, n):
        self._schema, self.n = schema, n

    def schema(self):
        return self._schema

    def output_partitioning(self):
        import ballista_b200 as bb

        return bb.Partitioning.UnknownPartitioning(self.n)

    def execute(self, p):
        return iter(())


def test_shuffle_writer_answers(bb):
    schema = pa.schema([("a", pa.uint32()), ("b", pa.string())])
    leaf = Leaf(schema, 7)
    w = bb.GpuShuffleWriterExec.try_new("job", 3, [0, 1], leaf, "/work", bb.Partitioning.Hash([0], 2), None)
    # schema() = the input's; output_partitioning() = the INPUT's partitioning, not the shuffle's (:483-493)
    assert w.schema() is schema and w.output_partitioning().partition_count() == 7
    assert w.shuffle_output_partitioning_().partition_count() == 2
    assert w.output_ordering() is None and w.children() == [leaf]
    assert str(w) == "GpuShuffleWriterExec: Hash([0], 2)"
    assert str(bb.GpuShuffleWriterExec.try_new("job", 3, [0], leaf, "/work", None, None)) == "GpuShuffleWriterExec"
    # with_new_children: the same writer over another child (:503-516)
    other = Leaf(pa.schema([("x", pa.int64())]), 3)
    w2 = w.with_new_children([other])
    assert w2 is not w and w2.children() == [other] and w2.schema() is other.schema()
    assert (w2.job_id, w2.stage_id, w2.partitions, w2.work_dir) == ("job", 3, [0, 1], "/work")
    assert w2.shuffle_output_partitioning is w.shuffle_output_partitioning and w2.output_partitioning().partition_count() == 3
    # with_partitions: what the task-group optimizer does to the stage's writer (:205-223, task_group.rs:60-64)
    w3 = w.with_partitions([4, 5, 6])
    assert w3.partitions == [4, 5, 6] and w.partitions == [0, 1] and w3.children() == [leaf] and w3.schema() is schema


def test_reader_and_coalesce_answers(bb):
    schema = pa.schema([("a", pa.int64())])
    reader = bb.GpuShuffleReaderExec([[], [], []], schema, None)
    assert reader.schema() is schema and reader.children() == [] and reader.output_ordering() is None
    assert reader.output_partitioning().partition_count() == 3 and reader.output_partitioning().kind == "unknown"
    with pytest.raises(bb.BgpuError, match="does not support with_new_children"):
        reader.with_new_children([])
    tasks = bb.CoalesceTasksExec(reader, [0, 2], schema)
    assert tasks.children() == [reader] and tasks.output_partitioning().partition_count() == 1
    other = bb.GpuShuffleReaderExec([[]], schema, None)
    t2 = tasks.with_new_children([other])
    assert t2.children() == [other] and t2.partitions == [0, 2] and tasks.children() == [reader]
    batches = bb.CoalesceBatchesExec(reader, None, 4096)
    assert batches.children() == [reader] and batches.schema() is schema
    assert batches.output_partitioning().partition_count() == 3
    b2 = batches.with_new_children([other])
    assert b2.children() == [other] and b2.target_batch_size == 4096


def test_aggregate_node_answers(bb):
    schema = pa.schema([("k", pa.int64()), ("v", pa.int64())])
    leaf = Leaf(schema, 5)
    node = bb.GpuAggregateExecNode(bb.AggregateMode.Partial, [0], [(bb.AggregateFunction.Sum, bb.expr.col(1), "s")], schema,
                                   inputs=[leaf])
    assert node.children() == [leaf] and node.output_partitioning().partition_count() == 5
    other = Leaf(schema, 2)
    n2 = node.with_new_children([other])
    assert n2.children() == [other] and node.children() == [leaf] and n2.group_by == [0] and n2.mode == node.mode

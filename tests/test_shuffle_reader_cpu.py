"""Host logic of the consumer-side operator mirrors (no GPU): CoalesceTasksExec fan-in
(/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:130-230), the buffering rules of CoalesceBatchesExec,
PartitionLocation bookkeeping and the FetchFailed mapping of GpuShuffleReaderExec
(/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:743-785)."""
import pyarrow as pa
import pytest


class FakeBatch:
    def __init__(self, n):
        self.num_rows = n


def test_coalesce_tasks_single_partition_passes_through(bb):
    parts = {3: ["x0", "x1"]}
    op = bb.CoalesceTasksExec(lambda p: parts[p], [3])
    assert list(op.execute()) == ["x0", "x1"]
    assert op.output_partitioning().partition_count() == 1


def test_coalesce_tasks_interleaves_and_drains_every_partition(bb):
    parts = {0: ["a0", "a1", "a2"], 1: ["b0"], 2: [], 3: ["d0", "d1"]}
    calls = []

    def execute(p):
        calls.append(p)
        return iter(parts[p])
    got = list(bb.CoalesceTasksExec(execute, [0, 1, 2, 3]).execute())
    assert calls == [0, 1, 2, 3]                      # every input partition is executed once
    assert sorted(got) == sorted(sum(parts.values(), []))
    assert got == ["a0", "b0", "d0", "a1", "d1", "a2"]  # one batch from each live partition in turn
    for p in parts:                                    # order inside a partition is kept
        assert [x for x in got if x in parts[p]] == parts[p]


def test_coalesce_batches_buffering_rules_without_concat(bb):
    # a batch of >= target rows arriving at an empty buffer passes through untouched; empty batches vanish; a lone
    # small batch is flushed at the end as it is
    big, small, empty = FakeBatch(200), FakeBatch(5), FakeBatch(0)
    out = list(bb.CoalesceBatchesExec([big, empty, small], None, target_batch_size=100))
    assert out == [big, small]
    assert list(bb.CoalesceBatchesExec([], None)) == []
    assert list(bb.CoalesceBatchesExec(lambda: [big], None, target_batch_size=100)) == [big]


def test_partition_location_from_shuffle_write(bb):
    w = bb.ShuffleWritePartition([4, 5], 7, "/work/job/2/7/u.arrow", 3, 120, 4096)
    loc = bb.PartitionLocation.from_shuffle_write("job", 2, w, bb.ExecutorMetadata(id="e1", host="h", port=50051))
    assert (loc.map_partitions, loc.job_id, loc.stage_id, loc.output_partition, loc.path) == ([4, 5], "job", 2, 7, w.path)
    assert (loc.partition_stats.num_rows, loc.partition_stats.num_batches, loc.partition_stats.num_bytes) == (120, 3, 4096)
    assert loc.executor_meta.id == "e1"


def test_reader_answers_and_fetch_failed_without_a_device(bb, tmp_path):
    schema = pa.schema([("a", pa.int64())])
    meta = bb.ExecutorMetadata(id="exec-7")
    missing = bb.PartitionLocation([3, 4], "job", 2, 0, str(tmp_path / "missing.arrow"), meta)
    reader = bb.GpuShuffleReaderExec([[missing], []], schema, None)
    assert reader.schema().equals(schema) and reader.output_partitioning().partition_count() == 2 and reader.children() == []
    with pytest.raises(bb.FetchFailed) as e:
        list(reader.execute(0))
    assert (e.value.executor_id, e.value.map_stage_id, e.value.map_partitions) == ("exec-7", 2, [3, 4])
    assert "Failed to open partition file" in e.value.message and "map_stage 2" in str(e.value)
    assert list(reader.execute(1)) == []  # no locations: an empty stream (shuffle_reader.rs:258-268)
    # locations whose statistics say "no rows" are never fetched (:248-252)
    empty = bb.PartitionLocation([0], "job", 2, 0, str(tmp_path / "missing.arrow"), meta, bb.PartitionStats(0, 0, 0))
    assert list(bb.GpuShuffleReaderExec([[empty]], schema, None).execute(0)) == []


def test_stats_for_partitions_as_the_reference_tests_them(bb):
    """test_stats_for_partitions_{empty,full,missing}
    (/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:876-940)."""
    S = bb.PartitionStats
    assert bb.stats_for_partitions(0, []) == {"num_rows": 0, "total_byte_size": 0, "column_statistics": []}
    full = [S(num_rows=10, num_batches=1, num_bytes=84), S(num_rows=4, num_batches=None, num_bytes=65)]
    assert bb.stats_for_partitions(0, full) == {"num_rows": 14, "total_byte_size": 149, "column_statistics": []}
    missing = [S(num_rows=10, num_batches=1, num_bytes=84), S(None, None, None)]
    assert bb.stats_for_partitions(0, missing) == {"num_rows": None, "total_byte_size": None, "column_statistics": []}
    # through the operator: every location of every reduce partition counts (shuffle_reader.rs:304-312)
    schema = pa.schema([("a", pa.int64()), ("b", pa.string())])
    locs = [[bb.PartitionLocation([0], "j", 1, 0, "p0", partition_stats=full[0])],
            [bb.PartitionLocation([0], "j", 1, 1, "p1", partition_stats=full[1])]]
    assert bb.GpuShuffleReaderExec(locs, schema, None).statistics() == {"num_rows": 14, "total_byte_size": 149,
                                                                       "column_statistics": [None, None]}

"""Host mirrors of the operators either side of the hot path on the CONSUMER side, with
the data staying on the B200:

  * `GpuShuffleReaderExec`  - `ShuffleReaderExec` (/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs):
    reduce partition p reads its file from every map task (`fetch_partition_local`, :743-785) and yields the batches in
    arrival order; here the files are decoded straight into device batches by libbgpu's IPC reader (`ipc.StreamReader`:
    LZ4_FRAME blocks decoded by a kernel). Inside one NVSwitch node `exchange.Exchange` bypasses the files altogether.
  * `CoalesceBatchesExec`   - DataFusion's operator the planner puts above every reader with target 8192 rows
    (/root/reference/ballista/scheduler/src/planner.rs:386-387, 409-413): small batches are concatenated on the device
    (`bgpu_batch_concat`).
  * `CoalesceTasksExec`     - fan-in of a task group's input partitions into one stream
    (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:130-230): one partition passes through
    (:144-146), several are interleaved (the reference's mpsc channel gives an arbitrary interleave; this mirror takes
    one batch from each live partition in turn, which is one of them).
"""
import random
from dataclasses import dataclass, field

from . import ipc
from ._lib import BGPU_EINVAL, BgpuError
from .partitioner import Partitioning
from .runtime import DeviceBatch


@dataclass
class ExecutorMetadata:
    """ballista.proto ExecutorMetadata (/root/reference/ballista/core/proto/ballista.proto:321-327), the fields the reader uses."""
    id: str = "local"
    host: str = "localhost"
    port: int = 0


@dataclass
class PartitionStats:
    num_rows: int = None
    num_batches: int = None
    num_bytes: int = None


@dataclass
class PartitionLocation:
    """ballista.proto PartitionLocation (/root/reference/ballista/core/proto/ballista.proto:260-270)."""
    map_partitions: list
    job_id: str
    stage_id: int
    output_partition: int
    path: str
    executor_meta: ExecutorMetadata = field(default_factory=ExecutorMetadata)
    partition_stats: PartitionStats = field(default_factory=PartitionStats)

    @classmethod
    def from_shuffle_write(cls, job_id, stage_id, w, executor_meta=None):
        """What the scheduler builds from a task's ShuffleWritePartition (execution_stage / partition_to_location)."""
        return cls(list(w.partitions), str(job_id), int(stage_id), int(w.output_partition), w.path,
                   executor_meta or ExecutorMetadata(), PartitionStats(w.num_rows, w.num_batches, w.num_bytes))


class FetchFailed(Exception):
    """BallistaError::FetchFailed(executor_id, map_stage_id, map_partitions, message): the one error with scheduler-visible
    semantics - the map stage is re-run (shuffle_reader.rs:751-759)."""

    def __init__(self, executor_id, map_stage_id, map_partitions, message):
        super().__init__(f"Shuffle fetch partition error from Executor {executor_id}, map_stage {map_stage_id}, "
                         f"map_partitions {map_partitions}: {message}")
        self.executor_id, self.map_stage_id, self.map_partitions, self.message = executor_id, map_stage_id, map_partitions, message


def stats_for_partitions(num_fields, partition_stats):
    """shuffle_reader.rs:315-337: exact row and byte totals of the partitions to read; one unknown statistic makes the
    total unknown (None = Precision::Absent); no column statistics."""
    num_rows, total_byte_size = 0, 0
    for part in partition_stats:
        num_rows = None if num_rows is None or part.num_rows is None else num_rows + part.num_rows
        total_byte_size = None if total_byte_size is None or part.num_bytes is None else total_byte_size + part.num_bytes
    return {"num_rows": num_rows, "total_byte_size": total_byte_size, "column_statistics": [None] * num_fields}


class GpuShuffleReaderExec:
    """`partition[p]` = the PartitionLocations reduce partition p reads; `execute(p)` yields DeviceBatches."""

    def __init__(self, partition, schema, ctx, stage_id=0):
        self.partition, self._schema, self.ctx, self.stage_id = [list(x) for x in partition], schema, ctx, stage_id
        self.metrics = {"num_rows": 0, "num_batches": 0, "bytes_read": 0}

    def schema(self):
        return self._schema

    def output_partitioning(self):
        # shuffle_reader.rs:203-207
        return Partitioning.UnknownPartitioning(len(self.partition))

    def children(self):
        return []

    def output_ordering(self):
        return None

    def statistics(self):
        # shuffle_reader.rs:304-312
        n = len(self._schema) if self._schema is not None else 0
        return stats_for_partitions(n, (loc.partition_stats for locs in self.partition for loc in locs))

    def with_new_children(self, _children):
        # shuffle_reader.rs:217-224
        raise BgpuError(BGPU_EINVAL, "Ballista ShuffleReaderExec does not support with_new_children()")

    def __str__(self):
        return f"GpuShuffleReaderExec: partitions={len(self.partition)}"

    def _fetch_partition_local(self, loc):
        try:
            with open(loc.path, "rb") as f:
                data = f.read()
        except OSError as e:
            raise FetchFailed(loc.executor_meta.id, loc.stage_id, loc.map_partitions,
                              f"Failed to open partition file at {loc.path}: {e!r}") from e
        try:
            return ipc.StreamReader(self.ctx, data), len(data)
        except BgpuError as e:
            raise FetchFailed(loc.executor_meta.id, loc.stage_id, loc.map_partitions,
                              f"Failed to new arrow StreamReader at {loc.path}: {e}") from e

    def execute(self, partition, shuffle=True):
        # locations whose statistics say "no rows" are skipped (shuffle_reader.rs:248-252); the rest are fetched in a
        # random order to spread the load over the executors (:270-271)
        locs = [l for l in self.partition[partition] if l.partition_stats.num_rows is None or l.partition_stats.num_rows > 0]
        if shuffle:
            random.shuffle(locs)
        for loc in locs:
            reader, nbytes = self._fetch_partition_local(loc)
            self.metrics["bytes_read"] += nbytes
            for b in reader:
                self.metrics["num_rows"] += b.num_rows
                self.metrics["num_batches"] += 1
                yield b
            reader.close()


class CoalesceBatchesExec:
    """DataFusion's CoalesceBatchesExec over DeviceBatches: a batch of >= target rows arriving at an empty buffer passes
    through untouched; otherwise batches are buffered until they hold >= target rows and concatenated; the rest is
    flushed at the end of the input."""

    def __init__(self, input_stream, ctx, target_batch_size=8192):
        self.input, self.ctx, self.target_batch_size = input_stream, ctx, int(target_batch_size)

    def children(self):
        return [self.input] if callable(getattr(self.input, "execute", None)) else []

    def with_new_children(self, children):
        return CoalesceBatchesExec(children[0], self.ctx, self.target_batch_size)

    def output_partitioning(self):
        child = getattr(self.input, "output_partitioning", None)
        return child() if callable(child) else Partitioning.UnknownPartitioning(1)

    def schema(self):
        return self.input.schema() if callable(getattr(self.input, "schema", None)) else None

    def execute(self, partition):
        """As a plan node over a child with `execute(p)` (the reader): partition p's batches, coalesced."""
        return iter(CoalesceBatchesExec(self.input.execute(partition), self.ctx, self.target_batch_size))

    def __iter__(self):
        buf, rows = [], 0
        for b in (self.input() if callable(self.input) else self.input):
            n = b.num_rows
            if n >= self.target_batch_size and not buf:
                yield b
                continue
            if n == 0:
                continue
            buf.append(b)
            rows += n
            if rows >= self.target_batch_size:
                yield buf[0] if len(buf) == 1 else DeviceBatch.concat(self.ctx, buf)
                buf, rows = [], 0
        if buf:
            yield buf[0] if len(buf) == 1 else DeviceBatch.concat(self.ctx, buf)


class CoalesceTasksExec:
    """`input` is the child: a plan node with `execute(p)` (kept as `.input`, what `children()` returns), or just that
    function (returns an iterable of batches); `partitions` the task group's input partitions."""

    def __init__(self, input, partitions, schema=None, order_by=None, ctx=None):
        """order_by: [(column index, descending, nulls_first)] when the child is ordered (coalesce_tasks.rs:148-173)."""
        self.input = input if callable(getattr(input, "execute", None)) else None
        self.execute_partition = input.execute if self.input is not None else input
        self.partitions, self._schema = list(partitions), schema
        self.order_by, self.ctx = (list(order_by) if order_by else None), ctx

    def children(self):
        return [self.input] if self.input is not None else []

    def with_new_children(self, children):
        # coalesce_tasks.rs:118-128: same partitions, new input
        return CoalesceTasksExec(children[0], self.partitions, self._schema, self.order_by, self.ctx)

    def output_ordering(self):
        return self.order_by  # coalesce_tasks.rs:109-111: the child's ordering survives the ordered fan-in

    def schema(self):
        return self._schema

    def output_partitioning(self):
        # coalesce_tasks.rs:105-107: one output partition
        return Partitioning.UnknownPartitioning(1)

    def execute(self, _partition=0):
        if len(self.partitions) == 1:
            yield from self.execute_partition(self.partitions[0])
            return
        if self.order_by:
            # the ordered fan-in: the reference merges the partitions' sorted streams (streaming_merge, ties to the lower
            # stream index); here: the batches concatenated in stream order, then one STABLE sort on the device
            batches = []
            for p in self.partitions:
                for b in self.execute_partition(p):
                    if not isinstance(b, DeviceBatch):
                        b = DeviceBatch.upload(self.ctx, b)
                    batches.append(b)
            if not batches:
                return
            ctx = self.ctx or batches[0].ctx
            whole = batches[0] if len(batches) == 1 else DeviceBatch.concat(ctx, batches)
            yield whole.sort([c for c, _, _ in self.order_by], [d for _, d, _ in self.order_by], [n for _, _, n in self.order_by])
            return
        live = [iter(self.execute_partition(p)) for p in self.partitions]
        while live:
            nxt = []
            for it in live:
                try:
                    b = next(it)
                except StopIteration:
                    continue
                nxt.append(it)
                yield b
            live = nxt

    __iter__ = execute

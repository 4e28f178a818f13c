"""Host mirror of Ballista's `ShuffleWriterExec` with the hash repartition on the GPU
(/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs).

`GpuShuffleWriterExec` keeps the reference operator's contract:
  * `try_new(job_id, stage_id, partitions, plan, work_dir, shuffle_output_partitioning)` (:165-190)
  * `schema()` = the input's schema, `output_partitioning()` = the INPUT partitioning (:483-493)
  * `execute_shuffle_write()` (:234-257, 328-449): drives the child, partitions every batch with
    `BatchPartitioner` (here: libbgpu's create_hashes + histogram/scan/scatter kernels), lazily opens one Arrow IPC
    stream writer (LZ4_FRAME) per non-empty output partition at `<work_dir>/<job_id>/<stage_id>/<p>/<uuid>.arrow`
    (:351-385), counts batches/rows, and reports `num_bytes` = the file length taken BEFORE `finish()` (:396-399);
    `None` partitioning writes one `data.arrow` (:258-326); anything else is
    "Invalid shuffle partitioning scheme" (:451-453)
  * `execute()` (:518-579) wraps the result as one batch with `result_schema()` (:590-597)
  * metrics `write_time`, `repart_time`, `input_rows`, `output_rows` (:139-163)

IPC framing: with `ipc="device"` (the default whenever the schema has no dictionary column) the
RecordBatch messages are produced by libbgpu (`ipc.py`: bodies laid out by one kernel, LZ4_FRAME-compressed per buffer
by another as the reference's writer compresses them - `compression=None` leaves them uncompressed, which the
reference's `StreamReader` reads just as well - one device-to-host copy, hand-built flatbuffer metadata) and appended
to the partition files as they are. With `ipc="host"` (and for dictionary columns)
pyarrow's writer stands in for arrow-ipc's `StreamWriter`, LZ4_FRAME as the reference writes. Inside one NVSwitch node
the files are bypassed altogether: see `exchange.py`.
"""
import os
import time
import uuid
from dataclasses import dataclass, field

import pyarrow as pa

from . import ipc
from ._lib import BGPU_EINVAL, BgpuError
from .partitioner import BatchPartitioner, Partitioning
from .runtime import Context, DeviceBatch


@dataclass
class ShuffleWritePartition:
    """ballista.proto ShuffleWritePartition (/root/reference/ballista/core/proto/ballista.proto:443-450)."""
    partitions: list
    output_partition: int
    path: str
    num_batches: int
    num_rows: int
    num_bytes: int


@dataclass
class ShuffleWriteMetrics:
    write_time: float = 0.0
    repart_time: list = field(default_factory=lambda: [0.0])  # shared with BatchPartitioner as its timer
    input_rows: int = 0
    output_rows: int = 0


def result_schema():
    """shuffle_writer.rs:590-597 + PartitionStats::arrow_struct_repr (serde/scheduler/mod.rs:132-146)."""
    stats = pa.struct([pa.field("num_rows", pa.uint64(), False), pa.field("num_batches", pa.uint64(), False),
                       pa.field("num_bytes", pa.uint64(), False)])
    return pa.schema([pa.field("partition", pa.uint32(), False), pa.field("path", pa.string(), False),
                      pa.field("partition_stats", stats, False)])


class _WriteTracker:
    def __init__(self, path, schema):
        self.path, self.num_batches, self.num_rows = path, 0, 0
        self._sink = open(path, "wb")
        self._writer = pa.ipc.new_stream(self._sink, schema, options=pa.ipc.IpcWriteOptions(compression="lz4"))

    def write(self, batch):
        self._writer.write_batch(batch)
        self.num_batches += 1
        self.num_rows += batch.num_rows

    def finish(self):
        """Returns the file length as the reference takes it: before the end-of-stream marker is written."""
        self._sink.flush()
        num_bytes = os.path.getsize(self.path)
        self._writer.close()
        self._sink.close()
        return num_bytes


class _IpcStreamFile:
    """A partition file fed with ready-made IPC messages (ipc.encode_partitions)."""

    def __init__(self, path, schema_msg):
        self.path, self.num_batches, self.num_rows = path, 0, 0
        self._sink = open(path, "wb")
        self._sink.write(schema_msg)

    def write_message(self, msg, rows):
        self._sink.write(msg)
        self.num_batches += 1
        self.num_rows += rows

    def finish(self):
        """Returns the file length as the reference takes it: before the end-of-stream marker is written."""
        self._sink.flush()
        num_bytes = os.path.getsize(self.path)
        self._sink.write(ipc.EOS)
        self._sink.close()
        return num_bytes


class GpuShuffleWriterExec:
    """ShuffleWriterExec whose `Partitioning::Hash` arm runs on the B200.

    `plan` is the child: any iterable (or zero-argument callable returning an iterable) of pyarrow RecordBatches with
    a `.schema`; `input_partitioning` is what `output_partitioning()` reports, as the reference reports its input's.
    """

    def __init__(self, job_id, stage_id, partitions, plan, work_dir, shuffle_output_partitioning, ctx=None,
                 input_schema=None, input_partitioning=None, ipc=None, compression="lz4"):
        self.ipc = ipc  # "device" (None = device) or "host" (pyarrow's writer: tests compare the two)
        self.compression = compression  # device framing: "lz4" (LZ4_FRAME, as the reference writes) or None
        self.job_id, self.stage_id, self.partitions = str(job_id), int(stage_id), list(partitions)
        self.plan, self.work_dir = plan, str(work_dir)
        self.shuffle_output_partitioning = shuffle_output_partitioning
        if input_schema is None:
            input_schema = getattr(plan, "schema", None)  # a stream's attribute, or a plan node's schema() method
            if callable(input_schema):
                input_schema = input_schema()
        self._schema = input_schema
        self._input_partitioning = input_partitioning or Partitioning.UnknownPartitioning(1)
        self.ctx = ctx
        self.metrics = ShuffleWriteMetrics()

    @classmethod
    def try_new(cls, job_id, stage_id, partitions, plan, work_dir, shuffle_output_partitioning, ctx=None, **kw):
        return cls(job_id, stage_id, partitions, plan, work_dir, shuffle_output_partitioning, ctx, **kw)

    # ---- ExecutionPlan answers -------------------------------------------------------------------------------
    def schema(self):
        return self._schema

    def output_partitioning(self):
        # the INPUT's partitioning (shuffle_writer.rs:483-493): the child's own answer when the child is a plan node
        child = getattr(self.plan, "output_partitioning", None)
        return child() if callable(child) else self._input_partitioning

    def output_ordering(self):
        return None  # shuffle_writer.rs:495-497

    def shuffle_output_partitioning_(self):
        return self.shuffle_output_partitioning

    def children(self):
        return [self.plan]

    def _copy(self, **changes):
        kw = dict(job_id=self.job_id, stage_id=self.stage_id, partitions=list(self.partitions), plan=self.plan,
                  work_dir=self.work_dir, shuffle_output_partitioning=self.shuffle_output_partitioning, ctx=self.ctx,
                  input_partitioning=self._input_partitioning, ipc=self.ipc, compression=self.compression)
        if "plan" not in changes:
            kw["input_schema"] = self._schema
        kw.update(changes)
        return GpuShuffleWriterExec.try_new(**kw)

    def with_new_children(self, children):
        """shuffle_writer.rs:503-516: the same writer over children[0]."""
        return self._copy(plan=children[0])

    def with_partitions(self, partitions):
        """shuffle_writer.rs:205-223 (used by the task-group optimizer): the same writer for another set of stage partitions."""
        return self._copy(partitions=[int(p) for p in partitions])

    def __str__(self):
        # DisplayAs (shuffle_writer.rs:459-476): "ShuffleWriterExec: <partitioning>" or just the name
        part = self.shuffle_output_partitioning
        if part is None:
            return "GpuShuffleWriterExec"
        return f"GpuShuffleWriterExec: {part.kind.capitalize()}({part.exprs}, {part.partition_count()})"

    def _stream(self):
        # the reference executes partition 0 of its child (shuffle_writer.rs:253: below a task group's
        # CoalesceTasksExec there is one); plain iterables and zero-argument callables stand in for a child stream
        p = self.plan
        if callable(getattr(p, "execute", None)):
            return iter(p.execute(0))
        return iter(p() if callable(p) else p)

    # ---- execute_shuffle_write ---------------------------------------------------------------------------------
    def execute_shuffle_write(self):
        base = os.path.join(self.work_dir, self.job_id, str(self.stage_id))
        shuffle_id = str(uuid.uuid4())
        m = self.metrics
        part = self.shuffle_output_partitioning
        if part is None:
            path = os.path.join(base, shuffle_id)
            os.makedirs(path, exist_ok=True)
            path = os.path.join(path, "data.arrow")
            w = None
            t0 = time.perf_counter()
            # write_stream_to_disk (utils.rs:179-219): zero-row batches are skipped, and num_bytes is the sum of the
            # batches' in-memory sizes (get_array_memory_size), not the file length (that is the hash arm's rule)
            mem_bytes = 0
            for batch in self._stream():
                if batch.num_rows == 0:
                    if self._schema is None:
                        self._schema = batch.schema
                    continue
                if isinstance(batch, DeviceBatch):
                    # a stage whose output is already on the device (a Final aggregate): framed by libbgpu like the hash
                    # arm's partitions, as one "partition" holding every row
                    schema = self._schema if self._schema is not None else batch.schema
                    if self.ipc == "host":
                        batch = batch.download()
                    else:
                        if w is None:
                            w = _IpcStreamFile(path, ipc.schema_message(schema, self.ctx))
                        mem_bytes += batch.memory_size
                        enc = ipc.encode_partitions(self.ctx, batch, [0, batch.num_rows], self.compression)
                        w.write_message(enc.message(0), batch.num_rows)
                        enc.close()
                        continue
                if w is None:
                    w = _WriteTracker(path, batch.schema)
                mem_bytes += batch.get_total_buffer_size()
                w.write(batch)
            if w is None:
                w = _WriteTracker(path, self._schema)
            w.finish()
            num_bytes = mem_bytes
            m.write_time += time.perf_counter() - t0
            m.input_rows += w.num_rows
            m.output_rows += w.num_rows
            return [ShuffleWritePartition([int(p) for p in self.partitions], 0, path, w.num_batches, w.num_rows, num_bytes)]
        if part.kind != "hash":
            raise BgpuError(BGPU_EINVAL, "Invalid shuffle partitioning scheme")
        n = part.partition_count()
        writers = [None] * n
        if self.ctx is None:
            self.ctx = Context(0)
        partitioner = BatchPartitioner.try_new(Partitioning.Hash(part.exprs, n), self.ctx, m.repart_time)
        mode = self.ipc
        if mode is None:
            sch = self._schema
            mode = "device"  # dictionary columns included: DictionaryBatch messages are framed on the device too
        if mode == "device":
            return self._write_device_ipc(base, shuffle_id, partitioner, n)

        def on_partition(output_partition, output_batch):
            # the partitioner never hands over an empty batch (shuffle_writer.rs:349)
            t0 = time.perf_counter()
            w = writers[output_partition]
            if w is None:
                d = os.path.join(base, str(output_partition))
                os.makedirs(d, exist_ok=True)
                w = writers[output_partition] = _WriteTracker(os.path.join(d, f"{shuffle_id}.arrow"), output_batch.schema)
            w.write(output_batch)
            m.output_rows += output_batch.num_rows
            m.write_time += time.perf_counter() - t0

        for input_batch in self._stream():
            m.input_rows += input_batch.num_rows
            partitioner.partition(input_batch, on_partition)
        part_locs = []
        for i, w in enumerate(writers):
            if w is not None:
                num_bytes = w.finish()
                part_locs.append(ShuffleWritePartition([int(p) for p in self.partitions], i, w.path, w.num_batches,
                                                       w.num_rows, num_bytes))
        return part_locs

    def _write_device_ipc(self, base, shuffle_id, partitioner, n):
        """Hash arm with the IPC messages framed by libbgpu: partition on the device, encode every non-empty partition
        of the batch in one call, append the messages to the partition files."""
        m = self.metrics
        writers = [None] * n
        schema_msg = None
        for input_batch in self._stream():
            m.input_rows += input_batch.num_rows
            if input_batch.num_rows == 0:
                continue
            if schema_msg is None:
                schema_msg = ipc.schema_message(self._schema if isinstance(input_batch, DeviceBatch) and self._schema is not None
                                                else input_batch.schema, self.ctx)
            dbatch = input_batch if isinstance(input_batch, DeviceBatch) else DeviceBatch.upload(self.ctx, input_batch)
            out, offs = partitioner.partition_device(dbatch)
            t0 = time.perf_counter()
            enc = ipc.encode_partitions(self.ctx, out, offs, self.compression)
            for p in range(n):
                rows = offs[p + 1] - offs[p]
                if rows == 0:
                    continue  # the writer never sees an empty batch (shuffle_writer.rs:349)
                w = writers[p]
                if w is None:
                    d = os.path.join(base, str(p))
                    os.makedirs(d, exist_ok=True)
                    w = writers[p] = _IpcStreamFile(os.path.join(d, f"{shuffle_id}.arrow"), schema_msg)
                w.write_message(enc.message(p), rows)
                m.output_rows += rows
            enc.close()
            m.write_time += time.perf_counter() - t0
        part_locs = []
        for i, w in enumerate(writers):
            if w is not None:
                num_bytes = w.finish()
                part_locs.append(ShuffleWritePartition([int(p) for p in self.partitions], i, w.path, w.num_batches,
                                                       w.num_rows, num_bytes))
        return part_locs

    # ---- execute: the metadata batch -------------------------------------------------------------------------------
    def execute(self, _partition=0):
        locs = self.execute_shuffle_write()
        stats = pa.StructArray.from_arrays(
            [pa.array([l.num_rows for l in locs], type=pa.uint64()), pa.array([l.num_batches for l in locs], type=pa.uint64()),
             pa.array([l.num_bytes for l in locs], type=pa.uint64())],
            fields=list(result_schema().field("partition_stats").type))
        batch = pa.RecordBatch.from_arrays(
            [pa.array([l.output_partition for l in locs], type=pa.uint32()), pa.array([l.path for l in locs], type=pa.string()),
             stats], schema=result_schema())
        return [batch]

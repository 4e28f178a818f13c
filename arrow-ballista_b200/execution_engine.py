"""Host mirror of the executor's stage-execution hook (SURVEY.md §8b "Stage execution hook"):
`ExecutionEngine::create_query_stage_exec(job_id, stage_id, plan, work_dir)` and `QueryStageExecutor`
(/root/reference/ballista/executor/src/execution_engine.rs:34-60), the swap point that needs no scheduler change.

  * `DefaultExecutionEngine` / `DefaultQueryStageExec` (:63-140): the root of every stage plan is a shuffle writer; it
    is re-created with the executor's own `work_dir` (the codec decodes it with "", serde/mod.rs:191) and
    `execute_query_stage` is its `execute_shuffle_write`. Anything else is
    "Plan passed to new_query_stage_exec is not a ShuffleWriterExec" (:88-92).
  * `GpuExecutionEngine` inspects the decoded plan. A map stage of the hot path has the shape the task-group optimizer
    leaves behind (/root/reference/ballista/core/src/physical_optimizer/task_group.rs:56-110: the CoalesceTasksExec
    inserted above the leaf is pulled up over every mapping node, Partial aggregates included):

        ShuffleWriterExec(Hash(keys, n)) -> [CoalesceTasksExec(partitions)] -> AggregateExec(Partial) -> source

    and becomes a `GpuQueryStageExec`: the source's batches are aggregated on the device (bgpu_agg_*), the state batch
    never leaves HBM on its way into the hash repartition (bgpu_hash_partition) and the IPC encoder
    (bgpu_ipc_encode_partitions); the partition files and the returned `ShuffleWritePartition`s are what the reference's
    stage would have produced. Every other plan goes to the `fallback` engine the embedder passes (in a Ballista
    executor: the reference's own DefaultExecutionEngine over DataFusion's CPU operators); without one, a stage whose
    root is a shuffle writer still runs its repartition on the device over whatever stream its child is, and there is
    NO CPU implementation of the aggregate in this package to fall back to.
"""
import time

from ._lib import BGPU_EINVAL, BgpuError
from .aggregate import AggregateFunction, AggregateMode
from .codec import GpuAggregateExecNode
from .runtime import Context, DeviceBatch
from .shuffle_reader import CoalesceTasksExec
from .shuffle_writer import GpuShuffleWriterExec

_NOT_A_WRITER = "Plan passed to new_query_stage_exec is not a ShuffleWriterExec"


class QueryStageExecutor:
    """execution_engine.rs:49-60."""

    def execute_query_stage(self, input_partitions, ctx=None):
        raise NotImplementedError

    def collect_plan_metrics(self):
        raise NotImplementedError

    def schema(self):
        raise NotImplementedError


def _recreate(writer, job_id, stage_id, work_dir, plan=None, ctx=None, input_schema=None):
    """execution_engine.rs:77-87: the same writer, with the executor's working directory."""
    return GpuShuffleWriterExec.try_new(job_id, stage_id, list(writer.partitions), writer.plan if plan is None else plan,
                                        work_dir, writer.shuffle_output_partitioning, ctx if ctx is not None else writer.ctx,
                                        input_schema=input_schema if input_schema is not None else writer.schema(),
                                        input_partitioning=writer.output_partitioning(), ipc=writer.ipc,
                                        compression=writer.compression)


class DefaultQueryStageExec(QueryStageExecutor):
    def __init__(self, shuffle_writer):
        self.shuffle_writer = shuffle_writer

    def execute_query_stage(self, input_partitions=None, ctx=None):
        if ctx is not None:
            self.shuffle_writer.ctx = ctx
        return self.shuffle_writer.execute_shuffle_write()

    def collect_plan_metrics(self):
        m = self.shuffle_writer.metrics
        return [{"write_time": m.write_time, "repart_time": m.repart_time[0], "input_rows": m.input_rows,
                 "output_rows": m.output_rows}]

    def schema(self):
        return self.shuffle_writer.schema()

    def __str__(self):
        return str(self.shuffle_writer)


class DefaultExecutionEngine:
    def create_query_stage_exec(self, job_id, stage_id, plan, work_dir, sender=None):
        if not isinstance(plan, GpuShuffleWriterExec):
            raise BgpuError(BGPU_EINVAL, _NOT_A_WRITER)
        return DefaultQueryStageExec(_recreate(plan, job_id, stage_id, work_dir))


def match_aggregate_stage(plan):
    """(coalesce | None, aggregate node, source) when `plan` is a stage of the hot path, else a string saying why not:

        map stage     ShuffleWriterExec(Hash(group columns, n)) -> [CoalesceTasksExec] -> AggregateExec(Partial) -> source
        reduce stage  ShuffleWriterExec(None | Hash)            -> [CoalesceTasksExec] -> AggregateExec(FinalPartitioned)
                                                                   -> [CoalesceBatchesExec] -> ShuffleReaderExec

    Purely structural: what the library cannot run (a key or argument type it has no kernel for) is reported by
    bgpu_agg_create_expr as BGPU_EUNSUPPORTED when the stage executes."""
    part = plan.shuffle_output_partitioning
    if part is not None and part.kind != "hash":
        return "the stage's output partitioning is neither hash nor none"
    node, coalesce = plan.plan, None
    if isinstance(node, CoalesceTasksExec):
        coalesce, node = node, node.input
    if not isinstance(node, GpuAggregateExecNode):
        return "the stage's child is not an aggregate"
    if node.mode not in (AggregateMode.Partial, AggregateMode.FinalPartitioned):
        return "the aggregate is neither Partial nor FinalPartitioned"
    if not 1 <= len(node.group_by) <= 8:
        return "1..8 group columns"
    if any(f not in (AggregateFunction.Sum, AggregateFunction.Avg, AggregateFunction.Count) for f, _, _ in node.aggr_expr):
        return "an aggregate function other than SUM / AVG / COUNT"
    if len(node.inputs) != 1 or not callable(getattr(node.inputs[0], "execute", None)):
        return "the aggregate needs exactly one executable input"
    if node.mode == AggregateMode.Partial:
        if part is None:
            return "a Partial aggregate whose stage does not hash-repartition its output"
        if any(k >= len(node.group_by) for k in part.exprs):
            # the state batch is (group columns ++ state columns): a stage planned by planner.rs:398-404 hashes group columns
            return "the output partitioning is not over group columns"
    return coalesce, node, node.inputs[0]


match_partial_aggregate_stage = match_aggregate_stage


class GpuQueryStageExec(QueryStageExecutor):
    """A map (Partial) or reduce (FinalPartitioned) stage on the device. With `merge_partitions` (the default) the input partitions of the task group feed ONE
    Partial aggregate - the reference runs one per partition below its CoalesceTasksExec and interleaves their state
    batches; merging them earlier yields fewer state rows for the same FinalPartitioned result. `merge_partitions=False`
    keeps one state batch per input partition, as the reference emits them. In a reduce stage the partitions hold
    disjoint keys, so one merged FinalPartitioned table gives the union of the per-partition results. The reduce
    stage's source (GpuShuffleReaderExec, CoalesceBatchesExec above it) hands over device batches; its output is
    written from the device as well (`data.arrow` for an un-partitioned final stage, shuffle_writer.rs:258-326)."""

    def __init__(self, shuffle_writer, aggregate, source, partitions=None, ctx=None, merge_partitions=True):
        self.shuffle_writer, self.aggregate, self.source = shuffle_writer, aggregate, source
        self.partitions, self.ctx, self.merge_partitions = partitions, ctx, merge_partitions
        self._schema = None
        self.metrics = {"aggregate_time": 0.0, "aggregate_input_rows": 0, "aggregate_output_rows": 0, "hot_kernel_ms": 0.0,
                        "last_path": ""}

    def _states(self, ctx, partitions):
        groups = ([list(partitions)] if partitions else []) if self.merge_partitions else [[p] for p in partitions]
        m = self.metrics
        for group in groups:
            t0 = time.perf_counter()
            agg = self.aggregate.instantiate(ctx)
            try:
                fed = 0
                for p in group:
                    for b in self.source.execute(p):
                        m["aggregate_input_rows"] += b.num_rows
                        agg.update(b)
                        fed += 1
                if not fed:
                    continue  # no batch at all (a reduce partition no map task wrote to): no output batch either
                state = agg.emit()
                m["hot_kernel_ms"] += agg.hot_kernel_ms
                m["last_path"] = agg.last_path
            finally:
                agg.close()
            m["aggregate_output_rows"] += state.num_rows
            m["aggregate_time"] += time.perf_counter() - t0
            yield state

    def execute_query_stage(self, input_partitions=None, ctx=None):
        ctx = ctx or self.ctx or self.shuffle_writer.ctx
        if ctx is None:
            ctx = self.ctx = Context(0)
        partitions = self.partitions if self.partitions is not None else list(input_partitions or [])
        if hasattr(ctx, "enable_times"):
            ctx.enable_times(True)
            ctx.times(reset=True)
            self._timed_ctx = ctx
        states = list(self._states(ctx, partitions))
        if not states:
            if self.shuffle_writer.shuffle_output_partitioning is not None:
                return []  # hash arm: writers open lazily, no batch = no file (shuffle_writer.rs:351-362)
            # un-partitioned output: the reference still writes an (empty) data.arrow and returns one zero-row entry
            # (utils.rs:179-219): an aggregate that saw no input emits the zero-row batch that carries the schema
            agg = self.aggregate.instantiate(ctx)
            try:
                states = [agg.emit()]
            finally:
                agg.close()
        self._schema = states[0].schema
        w = self.shuffle_writer
        mode = w.ipc
        if mode is None:
            import pyarrow as pa

            mode = "device"
        if mode == "host":
            states = [s.download() if isinstance(s, DeviceBatch) else s for s in states]
        self.shuffle_writer = _recreate(w, w.job_id, w.stage_id, w.work_dir, plan=states, ctx=ctx, input_schema=self._schema)
        return self.shuffle_writer.execute_shuffle_write()

    def collect_plan_metrics(self):
        m = self.shuffle_writer.metrics
        gpu = {}
        if getattr(self, "_timed_ctx", None) is not None:  # h2d_time / kernel_time / d2h_time of the whole stage (seconds)
            gpu = self._timed_ctx.times()
        return [dict(self.metrics), {"write_time": m.write_time, "repart_time": m.repart_time[0], "input_rows": m.input_rows,
                                     "output_rows": m.output_rows, **gpu}]

    def schema(self):
        """The stage's output schema = the aggregate's (known once the stage has run, or from the decoded plan)."""
        return self._schema if self._schema is not None else self.shuffle_writer.schema()

    def __str__(self):
        part = self.shuffle_writer.shuffle_output_partitioning
        out = "None" if part is None else f"Hash({part.exprs}, {part.partition_count()})"
        mode = {AggregateMode.Partial: "Partial", AggregateMode.FinalPartitioned: "FinalPartitioned"}.get(self.aggregate.mode)
        return (f"GpuQueryStageExec: GpuShuffleWriterExec({out}) <- "
                f"GpuAggregateExec({mode}, group_by={self.aggregate.group_by}) <- partitions {self.partitions}")


class GpuExecutionEngine:
    """`ExecutionEngine` for executors with a B200: hot-path map stages run on the device, the rest goes to `fallback`
    (default: `DefaultExecutionEngine`, whose shuffle writer still repartitions on the device)."""

    def __init__(self, ctx=None, fallback=None, merge_partitions=True):
        self.ctx, self.fallback, self.merge_partitions = ctx, fallback or DefaultExecutionEngine(), merge_partitions
        self.last_decision = ""

    def create_query_stage_exec(self, job_id, stage_id, plan, work_dir, sender=None):
        if not isinstance(plan, GpuShuffleWriterExec):
            raise BgpuError(BGPU_EINVAL, _NOT_A_WRITER)
        m = match_aggregate_stage(plan)
        if isinstance(m, str):
            self.last_decision = f"fallback: {m}"
            return self.fallback.create_query_stage_exec(job_id, stage_id, plan, work_dir, sender)
        coalesce, node, source = m
        self.last_decision = "gpu"
        writer = _recreate(plan, job_id, stage_id, work_dir, ctx=self.ctx)
        return GpuQueryStageExec(writer, node, source, list(coalesce.partitions) if coalesce is not None else None, self.ctx,
                                 self.merge_partitions)

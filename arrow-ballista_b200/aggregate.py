"""Host mirror of DataFusion's `AggregateExec` for the modes on Ballista's hot path
(plan shape /root/reference/ballista/scheduler/src/planner.rs:376-427; AggregateMode / AggregateFunction
numbering /root/reference/ballista/core/proto/datafusion.proto:1405-1411, 631-636).

`GpuAggregateExec` is what the north star calls the operator; the arithmetic runs in libbgpu
(bgpu_agg_create / update / emit).
"""
import ctypes as C

import pyarrow as pa

from ._lib import ArrowSchema, lib
from .runtime import DeviceBatch, _export, _release


def export_batch(batch):
    """(ArrowArray, ArrowSchema) C-Data structs of a host RecordBatch; the caller releases them (`_release`) or keeps
    them for `GpuAggregateExec.push`."""
    return _export(batch)


class AggregateMode:
    Partial = 0
    Final = 1
    FinalPartitioned = 2
    Single = 3


class AggregateFunction:
    Min = 0
    Max = 1
    Sum = 2
    Avg = 3
    Count = 4

    @staticmethod
    def parse(fn):
        if isinstance(fn, int):
            return fn
        return {"min": 0, "max": 1, "sum": 2, "avg": 3, "count": 4}[fn.lower()]


class GpuAggregateExec:
    """AggregateExec(mode, group_by, aggr_expr, input, input_schema).

    group_by: column indices into `input_schema` (the raw input of the Partial aggregate).
    aggr_expr: [(fn, arg_col | None, name)]; None = COUNT(*).
    In Final modes the batches passed to `update` are Partial state batches (group columns first).
    """

    def __init__(self, ctx, mode, group_by, aggr_expr, input_schema, filter=None):
        """aggr_expr: [(fn, arg, name)], arg = None (COUNT(*)), a column index, or an `expr.Expr` tree evaluated
        inside the aggregate as DataFusion does; `filter` = an optional `expr.Expr` predicate fused below the
        aggregate (Q1's FilterExec, /root/reference/benchmarks/queries/q1.sql)."""
        from ._lib import ExprNode
        from .expr import Expr, col

        self.ctx, self.mode, self.input_schema, self.filter = ctx, mode, input_schema, filter
        self.group_by = list(group_by)
        self.aggr_expr = []
        nodes, offsets = [], [0]
        for f, c, n in aggr_expr:
            e = None if c is None else (c if hasattr(c, "to_nodes") else col(int(c)))
            self.aggr_expr.append((AggregateFunction.parse(f), e, str(n)))
            if e is not None:
                nodes.extend(e.to_nodes())
            offsets.append(len(nodes))
        fnodes = filter.to_nodes() if filter is not None else []
        s = ArrowSchema()
        input_schema._export_to_c(C.addressof(s))
        n = len(self.aggr_expr)
        gcols = (C.c_int32 * len(self.group_by))(*self.group_by)
        fns = (C.c_int32 * max(n, 1))(*[f for f, _, _ in self.aggr_expr])
        names = (C.c_char_p * max(n, 1))(*[nm.encode() for _, _, nm in self.aggr_expr])
        node_arr = (ExprNode * max(len(nodes), 1))(*nodes)
        off_arr = (C.c_int32 * len(offsets))(*offsets)
        filt_arr = (ExprNode * max(len(fnodes), 1))(*fnodes)
        self._h = C.c_void_p()
        try:
            ctx.check(lib().bgpu_agg_create_expr(ctx._h, mode, C.byref(s), gcols, len(self.group_by), fns, node_arr,
                                                 off_arr, names, n, filt_arr, len(fnodes), C.byref(self._h)))
        finally:
            _release(s)

    @property
    def last_path(self):
        """Which kernel path the last update took (introspection)."""
        return lib().bgpu_agg_last_path(self._h).decode()

    @property
    def hot_kernel_ms(self):
        """Device time (ms) of this aggregate's fused single-pass kernel launches so far (introspection)."""
        return float(lib().bgpu_agg_hot_kernel_ms(self._h))

    def update(self, batch):
        """update_batch / merge_batch with one batch (DeviceBatch, or a host RecordBatch that is uploaded)."""
        if not isinstance(batch, DeviceBatch):
            batch = DeviceBatch.upload(self.ctx, batch)
        self.ctx.check(lib().bgpu_agg_update(self._h, batch._h))

    def push(self, batch):
        """One HOST RecordBatch of the child stream through the streamed input (bgpu_agg_push_host): packed into pinned
        staging by the calling thread (thread-safe: one thread per input partition may push at once), uploaded and
        aggregated chunk-wise. `batch` is a pyarrow RecordBatch or an (ArrowArray, ArrowSchema | None) pair already
        exported with `export_batch` (what a Rust caller holds: FFI structs cost nothing to make)."""
        if isinstance(batch, tuple):
            a, s = batch
            rc = lib().bgpu_agg_push_host(self._h, C.byref(a), C.byref(s) if s is not None else None)
            self.ctx.check(rc)
            return
        a, s = export_batch(batch)
        try:
            self.ctx.check(lib().bgpu_agg_push_host(self._h, C.byref(a), C.byref(s)))
        finally:
            _release(a)
            _release(s)

    def flush(self):
        """Submit the partly filled staging chunk and wait for every launch's outcome (bgpu_agg_flush)."""
        self.ctx.check(lib().bgpu_agg_flush(self._h))

    @property
    def ingest_stats(self):
        out = (C.c_uint64 * 8)()
        self.ctx.check(lib().bgpu_agg_ingest_stats(self._h, out))
        keys = ("rows", "batches", "wire_bytes", "arrow_bytes", "chunks", "async_launches", "replays", "one_pass_columns")
        return dict(zip(keys, [int(x) for x in out]))

    def reset(self):
        """A new input stream over the same plan (`execute` of the next partition): bgpu_agg_reset."""
        self.ctx.check(lib().bgpu_agg_reset(self._h))

    def emit(self):
        """End of input: the state batch (Partial) or the final values. Returns a DeviceBatch."""
        out = C.c_void_p()
        self.ctx.check(lib().bgpu_agg_emit(self._h, C.byref(out)))
        return DeviceBatch(self.ctx, out)

    @property
    def num_groups(self):
        return int(lib().bgpu_agg_num_groups(self._h))

    def execute(self, batches):
        """Run over an iterable of batches and return the output as a host RecordBatch."""
        for b in batches:
            self.update(b)
        return self.emit().download()

    def close(self):
        if getattr(self, "_h", None):
            lib().bgpu_agg_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()


def debug_compile(mode, group_by, aggr_expr, input_schema, filter=None, flags=0, threads=256, rows_per_thread=4,
                  min_ctas=2, out_prefix=None):
    """Generate and NVRTC-compile the run-time specialised kernels of a plan without a device
    (bgpu_agg_debug_compile): used by the CPU test-suite and to inspect the SASS offline. Raises BgpuError."""
    from ._lib import BgpuError, ExprNode
    from .expr import col

    L = lib()
    h = C.c_void_p()
    L.bgpu_ctx_create(0, C.byref(h))  # fails with ENODEVICE on a CPU box; the handle still carries error text
    try:
        nodes, offsets, fns = [], [0], []
        for f, c, _ in aggr_expr:
            e = None if c is None else (c if hasattr(c, "to_nodes") else col(int(c)))
            fns.append(AggregateFunction.parse(f))
            if e is not None:
                nodes.extend(e.to_nodes())
            offsets.append(len(nodes))
        fnodes = filter.to_nodes() if filter is not None else []
        s = ArrowSchema()
        input_schema._export_to_c(C.addressof(s))
        try:
            rc = L.bgpu_agg_debug_compile(
                h, mode, C.byref(s), (C.c_int32 * len(group_by))(*group_by), len(group_by),
                (C.c_int32 * max(len(fns), 1))(*fns), (ExprNode * max(len(nodes), 1))(*nodes),
                (C.c_int32 * len(offsets))(*offsets), len(fns), (ExprNode * max(len(fnodes), 1))(*fnodes), len(fnodes),
                flags, threads, rows_per_thread, min_ctas, out_prefix.encode() if out_prefix else None)
        finally:
            _release(s)
        if rc != 0:
            raise BgpuError(rc, L.bgpu_last_error(h).decode(errors="replace"))
    finally:
        L.bgpu_ctx_destroy(h)

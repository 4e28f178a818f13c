"""Host mirror of DataFusion's `BatchPartitioner` / `Partitioning` as used by Ballista's ShuffleWriterExec
(/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-339 try_new, :346-391 partition).

Same names and argument meaning as the reference API; the work happens in libbgpu (bgpu_hash_partition).
"""
import ctypes as C
import time

import pyarrow as pa

from ._lib import BGPU_EINVAL, BgpuError, lib
from .runtime import Context, DeviceBatch


class Partitioning:
    """datafusion::physical_plan::Partitioning. Only the Hash arm is on the hot path."""

    def __init__(self, kind, exprs=None, n=0):
        self.kind, self.exprs, self.n = kind, list(exprs or []), int(n)

    @classmethod
    def Hash(cls, exprs, n):
        """exprs: column indices (the reference's exprs are `Column` physical exprs, shuffle_writer.rs:626)."""
        return cls("hash", exprs, n)

    @classmethod
    def RoundRobinBatch(cls, n):
        return cls("round_robin", None, n)

    @classmethod
    def UnknownPartitioning(cls, n):
        return cls("unknown", None, n)

    def partition_count(self):
        return self.n


class BatchPartitioner:
    """`BatchPartitioner::try_new(partitioning, timer)` + `.partition(batch, f)`.

    `timer` accumulates seconds spent repartitioning (the reference's `repart_time` metric,
    shuffle_writer.rs:139-150). The callback receives `(partition, RecordBatch)` in ascending partition order,
    empty partitions skipped (shuffle_writer.rs:349).
    """

    def __init__(self, partitioning, ctx, timer=None):
        self.partitioning, self.ctx, self.timer = partitioning, ctx, timer if timer is not None else [0.0]

    @classmethod
    def try_new(cls, partitioning, ctx=None, timer=None):
        if partitioning.kind != "hash":
            # reference: shuffle_writer.rs:451-453 "Invalid shuffle partitioning scheme"
            raise BgpuError(BGPU_EINVAL, f"Unsupported repartitioning scheme {partitioning.kind}")
        if partitioning.n <= 0 or not partitioning.exprs:
            raise BgpuError(BGPU_EINVAL, "Hash partitioning needs >=1 key expression and n > 0")
        return cls(partitioning, ctx if ctx is not None else Context(), timer)

    def partition_device(self, dbatch, want_hashes_ptr=None):
        """Device batch in -> (partitioned DeviceBatch, offsets list[n+1])."""
        n = self.partitioning.n
        keys = (C.c_int32 * len(self.partitioning.exprs))(*self.partitioning.exprs)
        offs = (C.c_uint64 * (n + 1))()
        out = C.c_void_p()
        t0 = time.perf_counter()
        self.ctx.check(lib().bgpu_hash_partition(self.ctx._h, dbatch._h, keys, len(keys), n, C.byref(out), offs,
                                                 C.c_void_p(want_hashes_ptr) if want_hashes_ptr else None))
        self.timer[0] += time.perf_counter() - t0
        return DeviceBatch(self.ctx, out), list(offs)

    def partition(self, batch, f):
        """Host RecordBatch in; calls f(partition, RecordBatch) per non-empty output partition."""
        from .runtime import _export, _release
        from ._lib import ArrowArray, ArrowSchema

        n = self.partitioning.n
        keys = (C.c_int32 * len(self.partitioning.exprs))(*self.partitioning.exprs)
        offs = (C.c_uint64 * (n + 1))()
        a, s = _export(batch)
        oa, os_ = ArrowArray(), ArrowSchema()
        t0 = time.perf_counter()
        try:
            self.ctx.check(lib().bgpu_hash_partition_host(self.ctx._h, C.byref(a), C.byref(s), keys, len(keys), n,
                                                          C.byref(oa), C.byref(os_), offs))
        finally:
            _release(a)
            _release(s)
        self.timer[0] += time.perf_counter() - t0
        whole = pa.RecordBatch._import_from_c(C.addressof(oa), C.addressof(os_))
        for p in range(n):
            lo, hi = int(offs[p]), int(offs[p + 1])
            if hi > lo:
                f(p, whole.slice(lo, hi - lo))

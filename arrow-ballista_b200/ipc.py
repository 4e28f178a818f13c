"""Arrow IPC stream framing of partitioned device batches (bgpu_ipc_*).

Host mirror of what `ShuffleWriterExec` does with arrow-ipc's `StreamWriter`
(/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:351-399): a partition file is the schema message,
one RecordBatch message per batch written to it, and the end-of-stream marker. The message bodies are laid out on the
device by one kernel and fetched with one copy; only the few hundred bytes of flatbuffer metadata are host work.
"""
import ctypes as C

from ._lib import ArrowSchema, BgpuError, lib

EOS = b"\xff\xff\xff\xff\x00\x00\x00\x00"


class _HostContext:
    """A bgpu_ctx that may have no device behind it: enough for the host-only entry points (error text lives here)."""

    def __init__(self):
        self._h = C.c_void_p()
        lib().bgpu_ctx_create(0, C.byref(self._h))  # ENODEVICE on a CPU box; the handle still carries error text

    def check(self, rc):
        if rc != 0:
            raise BgpuError(rc, lib().bgpu_last_error(self._h).decode())

    def __del__(self):
        if self._h:
            lib().bgpu_ctx_destroy(self._h)
            self._h = None


def _take(p, n):
    try:
        return C.string_at(p.value, n.value)
    finally:
        lib().bgpu_ipc_free(p, 0)


def schema_message(schema, ctx=None):
    """The encapsulated Schema message of a pyarrow schema (bytes). Host only."""
    ctx = ctx or _HostContext()
    s = ArrowSchema()
    schema._export_to_c(C.addressof(s))
    p, n = C.c_void_p(), C.c_uint64()
    try:
        ctx.check(lib().bgpu_ipc_schema_message(ctx._h, C.byref(s), C.byref(p), C.byref(n)))
    finally:
        if s.release:
            C.CFUNCTYPE(None, C.c_void_p)(s.release)(C.addressof(s))
    return _take(p, n)


def record_batch_metadata(rows, nodes, buffers, body_len, ctx=None):
    """Prefix + RecordBatch metadata of one message: nodes = [(length, null_count)], buffers = [(offset, length)]. Host only."""
    ctx = ctx or _HostContext()
    nv = (C.c_int64 * max(1, 2 * len(nodes)))(*[int(x) for pair in nodes for x in pair])
    bv = (C.c_int64 * max(1, 2 * len(buffers)))(*[int(x) for pair in buffers for x in pair])
    p, n = C.c_void_p(), C.c_uint64()
    ctx.check(lib().bgpu_ipc_record_batch_metadata(ctx._h, rows, nv, len(nodes), bv, len(buffers), body_len, C.byref(p), C.byref(n)))
    return _take(p, n)


class EncodedPartitions:
    """The RecordBatch messages of one partitioned batch, one per non-empty partition, in pinned host memory."""

    def __init__(self, ptr, msg_off):
        self._ptr, self.msg_off = ptr, msg_off

    def message(self, p):
        """memoryview over the message of partition p (empty for an empty partition)."""
        lo, hi = self.msg_off[p], self.msg_off[p + 1]
        if hi == lo or not self._ptr:
            return memoryview(b"")
        return memoryview((C.c_ubyte * (hi - lo)).from_address(self._ptr.value + lo))

    @property
    def total_bytes(self):
        return self.msg_off[-1]

    def close(self):
        if self._ptr:
            lib().bgpu_ipc_free(self._ptr, 1)
            self._ptr = None

    def __del__(self):
        self.close()


def encode_partitions(ctx, device_batch, offsets):
    """Rows [offsets[p], offsets[p + 1]) of `device_batch` for every p, each as one encapsulated RecordBatch message."""
    n = len(offsets) - 1
    offs = (C.c_uint64 * (n + 1))(*[int(o) for o in offsets])
    msg_off = (C.c_uint64 * (n + 1))()
    p = C.c_void_p()
    ctx.check(lib().bgpu_ipc_encode_partitions(ctx._h, device_batch._h, offs, n, C.byref(p), msg_off))
    return EncodedPartitions(p if p.value else None, list(msg_off))

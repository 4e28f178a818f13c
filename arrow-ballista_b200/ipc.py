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
            raise BgpuError(rc, lib().bgpu_last_error(self._h).decode(errors="replace"))

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


def record_batch_metadata(rows, nodes, buffers, body_len, ctx=None, compression=-1):
    """Prefix + RecordBatch metadata of one message: nodes = [(length, null_count)], buffers = [(offset, length)]; compression
    0 marks the buffers as LZ4_FRAME-compressed. Host only."""
    ctx = ctx or _HostContext()
    nv = (C.c_int64 * max(1, 2 * len(nodes)))(*[int(x) for pair in nodes for x in pair])
    bv = (C.c_int64 * max(1, 2 * len(buffers)))(*[int(x) for pair in buffers for x in pair])
    p, n = C.c_void_p(), C.c_uint64()
    ctx.check(lib().bgpu_ipc_record_batch_metadata(ctx._h, rows, nv, len(nodes), bv, len(buffers), body_len, compression, C.byref(p), C.byref(n)))
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


def encode_partitions(ctx, device_batch, offsets, compression=None):
    """Rows [offsets[p], offsets[p + 1]) of `device_batch` for every p, each as one encapsulated RecordBatch message;
    compression="lz4": buffers LZ4_FRAME-compressed on the device, as the reference's writer compresses them."""
    n = len(offsets) - 1
    offs = (C.c_uint64 * (n + 1))(*[int(o) for o in offsets])
    msg_off = (C.c_uint64 * (n + 1))()
    p = C.c_void_p()
    if compression not in (None, "lz4"):
        raise ValueError("compression must be None or 'lz4'")
    ctx.check(lib().bgpu_ipc_encode_partitions(ctx._h, device_batch._h, offs, n, 0 if compression == "lz4" else -1, C.byref(p), msg_off))
    return EncodedPartitions(p if p.value else None, list(msg_off))


class StreamReader:
    """Arrow IPC stream (bytes-like, host memory) -> DeviceBatches: the device-side `AsyncStreamReader`
    (/root/reference/ballista/core/src/async_reader/mod.rs:108-258). Reads the reference's LZ4_FRAME partition files
    (blocks decoded by a kernel) and the uncompressed ones `encode_partitions` writes. Iterable."""

    def __init__(self, ctx, data):
        import numpy as np

        self.ctx = ctx
        self._buf = np.frombuffer(data, dtype=np.uint8)  # zero-copy view; keeps `data` alive for the reader
        self._h = C.c_void_p()
        ctx.check(lib().bgpu_ipc_reader_open(ctx._h, C.c_void_p(self._buf.ctypes.data), self._buf.size, C.byref(self._h)))
        self.num_batches = self.num_rows = 0

    @property
    def schema(self):
        import pyarrow as pa

        s = ArrowSchema()
        self.ctx.check(lib().bgpu_ipc_reader_schema(self._h, C.byref(s)))
        return pa.Schema._import_from_c(C.addressof(s))

    def maybe_next(self):
        """The next batch on the device, or None at the end of the stream (`maybe_next`, async_reader/mod.rs:168)."""
        from .runtime import DeviceBatch

        out = C.c_void_p()
        self.ctx.check(lib().bgpu_ipc_reader_next(self._h, C.byref(out)))
        if not out.value:
            return None
        b = DeviceBatch(self.ctx, out)
        self.num_batches += 1
        self.num_rows += b.num_rows
        return b

    def __iter__(self):
        while True:
            b = self.maybe_next()
            if b is None:
                return
            yield b

    def close(self):
        if getattr(self, "_h", None):
            lib().bgpu_ipc_reader_close(self._h)
            self._h = None

    def __del__(self):
        self.close()


def lz4_frame_decode_host(frame, out_len, ctx=None):
    """TEST HOOK: the library's LZ4 frame walker + block decoder run on the host with one lane (the kernel's logic)."""
    ctx = ctx or _HostContext()
    out = (C.c_ubyte * max(1, out_len))()
    src = (C.c_ubyte * max(1, len(frame))).from_buffer_copy(bytes(frame) or b"\0")
    ctx.check(lib().bgpu_debug_lz4_frame_decode_host(ctx._h, src, len(frame), out, out_len))
    return bytes(out[:out_len])

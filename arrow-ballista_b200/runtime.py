"""Host-side plumbing above the C ABI: contexts and device-resident RecordBatches (pyarrow in, pyarrow out)."""
import ctypes as C

import numpy as np
import pyarrow as pa

from ._lib import ArrowArray, ArrowSchema, BgpuError, lib


class Context:
    """One per task: a CUDA stream + pool (bgpu_ctx). `stream` may be an existing cudaStream_t handle (int)."""

    def __init__(self, device=0, stream=None):
        self._h = C.c_void_p()
        L = lib()
        if stream is None:
            rc = L.bgpu_ctx_create(device, C.byref(self._h))
        else:
            rc = L.bgpu_ctx_create_on_stream(device, C.c_void_p(stream), C.byref(self._h))
        if rc != 0:
            msg = L.bgpu_last_error(self._h).decode(errors="replace") if self._h else "allocation failed"
            L.bgpu_ctx_destroy(self._h)
            self._h = None
            raise BgpuError(rc, msg)

    def check(self, rc):
        if rc != 0:
            raise BgpuError(rc, lib().bgpu_last_error(self._h).decode(errors="replace"))

    def synchronize(self):
        self.check(lib().bgpu_ctx_synchronize(self._h))

    @property
    def kernel_launches(self):
        return int(lib().bgpu_ctx_kernel_launches(self._h))

    @property
    def stream(self):
        return lib().bgpu_ctx_stream(self._h)

    def enable_times(self, on=True):
        """Account the GPU time of every call on this context as h2d / kernel / d2h (bgpu_ctx_enable_times)."""
        self.check(lib().bgpu_ctx_enable_times(self._h, 1 if on else 0))

    def times(self, reset=False):
        """{"h2d_time", "kernel_time", "d2h_time"} in seconds of GPU time so far (bgpu_ctx_times): the metrics a GPU
        operator reports next to the reference's write_time / repart_time (shuffle_writer.rs:139-163)."""
        out = (C.c_double * 3)()
        self.check(lib().bgpu_ctx_times(self._h, out, 1 if reset else 0))
        return {"h2d_time": out[0] * 1e-3, "kernel_time": out[1] * 1e-3, "d2h_time": out[2] * 1e-3}

    def close(self):
        if getattr(self, "_h", None):
            lib().bgpu_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def _export(batch):
    """pyarrow RecordBatch -> (ArrowArray, ArrowSchema) C structs owned by the caller (released on import/del)."""
    a, s = ArrowArray(), ArrowSchema()
    batch._export_to_c(C.addressof(a), C.addressof(s))
    return a, s


def _release(st):
    if st.release:
        C.CFUNCTYPE(None, C.c_void_p)(st.release)(C.addressof(st))


class DeviceBatch:
    """A RecordBatch in device memory (bgpu_batch)."""

    def __init__(self, ctx, handle, keepalive=None):
        self.ctx, self._h, self._keep = ctx, handle, keepalive

    @classmethod
    def upload(cls, ctx, batch):
        if isinstance(batch, pa.Table):
            batch = batch.combine_chunks().to_batches()[0] if batch.num_rows else pa.RecordBatch.from_pylist(
                [], schema=batch.schema)
        a, s = _export(batch)
        h = C.c_void_p()
        try:
            ctx.check(lib().bgpu_batch_upload(ctx._h, C.byref(a), C.byref(s), C.byref(h)))
        finally:
            _release(a)
            _release(s)
        return cls(ctx, h)

    @classmethod
    def wrap_device(cls, ctx, schema, columns, n_rows, keepalive=None):
        """Borrow device buffers. columns[i] = dict(values=ptr, validity=ptr|None, offsets=ptr|None,
        null_count=int, dictionary=dict(..., length=int)|None); pointers are ints (e.g. tensor.data_ptr())."""
        s = ArrowSchema()
        schema._export_to_c(C.addressof(s))
        keep = []

        def mk(col, length, fmt_is_var):
            arr = ArrowArray()
            bufs = [col.get("validity")]
            if fmt_is_var:
                bufs.append(col["offsets"])
            bufs.append(col["values"])
            cb = (C.c_void_p * len(bufs))(*[C.c_void_p(b) if b else None for b in bufs])
            keep.append(cb)
            arr.length, arr.null_count, arr.offset = length, col.get("null_count", 0), 0
            arr.n_buffers, arr.buffers = len(bufs), C.cast(cb, C.POINTER(C.c_void_p))
            arr.n_children, arr.children, arr.release = 0, None, None
            if col.get("dictionary") is not None:
                d = col["dictionary"]
                da = mk(d, d["length"], d.get("offsets") is not None)
                keep.append(da)
                arr.dictionary = C.pointer(da)
            return arr

        kids = []
        for f, col in zip(schema, columns):
            t = f.type.value_type if pa.types.is_dictionary(f.type) else f.type
            is_var = (pa.types.is_string(t) or pa.types.is_binary(t)) and not pa.types.is_dictionary(f.type)
            kids.append(mk(col, n_rows, is_var))
        kid_ptrs = (C.POINTER(ArrowArray) * len(kids))(*[C.pointer(k) for k in kids])
        top = ArrowArray()
        nullbuf = (C.c_void_p * 1)(None)
        top.length, top.null_count, top.offset = n_rows, 0, 0
        top.n_buffers, top.buffers = 1, C.cast(nullbuf, C.POINTER(C.c_void_p))
        top.n_children, top.children = len(kids), C.cast(kid_ptrs, C.POINTER(C.POINTER(ArrowArray)))
        h = C.c_void_p()
        try:
            ctx.check(lib().bgpu_batch_wrap_device(ctx._h, C.byref(top), C.byref(s), C.byref(h)))
        finally:
            _release(s)
        return cls(ctx, h, keepalive=(keepalive, keep, kids))

    @property
    def num_rows(self):
        return int(lib().bgpu_batch_num_rows(self._h))

    @property
    def memory_size(self):
        """Bytes of the batch's Arrow buffers in device memory (the reference's `get_array_memory_size` of a batch)."""
        return int(lib().bgpu_batch_memory_size(self._h))

    @property
    def num_columns(self):
        return int(lib().bgpu_batch_num_columns(self._h))

    @property
    def schema(self):
        """The batch's Arrow schema (a zero-row download: no column data moves)."""
        return self.download(0, 0).schema

    def download(self, offset=0, length=None):
        """Device -> host copy of rows [offset, offset+length) as a pyarrow RecordBatch."""
        if length is None:
            length = self.num_rows - offset
        a, s = ArrowArray(), ArrowSchema()
        self.ctx.check(lib().bgpu_batch_download(self.ctx._h, self._h, offset, length, C.byref(a), C.byref(s)))
        return pa.RecordBatch._import_from_c(C.addressof(a), C.addressof(s))

    def column_tensor(self, i):
        """A torch int64 tensor [rows, width / 8] holding a COPY of fixed-width column i's values (width 8 or 16), made
        on the ctx stream. Plumbing for the bench's and the tests' device-side verification, not part of the reference's
        interface."""
        import torch
        from cuda.bindings import runtime as cudart

        a, s = ArrowArray(), ArrowSchema()
        self.ctx.check(lib().bgpu_batch_export_device(self.ctx._h, self._h, C.byref(a), C.byref(s)))
        try:
            fmt = s.children[i].contents.format.decode()
            width = 16 if fmt.startswith("d:") else 8
            if not (fmt.startswith("d:") or fmt in ("l", "L", "g")):
                raise TypeError(f"column_tensor: 8- or 16-byte columns only, got {fmt}")
            child = a.children[i].contents
            n = int(child.length)
            out = torch.empty((n, width // 8), dtype=torch.int64, device=torch.device("cuda", torch.cuda.current_device()))
            if n:
                err, = cudart.cudaMemcpyAsync(out.data_ptr(), child.buffers[1], n * width,
                                              cudart.cudaMemcpyKind.cudaMemcpyDeviceToDevice, self.ctx.stream or 0)
                if int(err) != 0:
                    raise RuntimeError(f"cudaMemcpyAsync: {err}")
                self.ctx.synchronize()
            return out
        finally:
            _release(a)
            _release(s)

    @classmethod
    def concat(cls, ctx, batches):
        """Rows of `batches` (one schema) in order as one DeviceBatch (bgpu_batch_concat; `concat_batches` of
        CoalesceBatchesExec, planned at /root/reference/ballista/scheduler/src/planner.rs:386-387)."""
        hs = (C.c_void_p * len(batches))(*[b._h for b in batches])
        out = C.c_void_p()
        ctx.check(lib().bgpu_batch_concat(ctx._h, hs, len(batches), C.byref(out)))
        return cls(ctx, out)

    def sort(self, sort_cols, descending=None, nulls_first=None):
        """A stable sort of the rows by the columns `sort_cols` (bgpu_batch_sort); `descending` / `nulls_first`: one flag
        per key (DataFusion's SortOptions; default ascending, nulls last)."""
        k = len(sort_cols)
        cols = (C.c_int32 * k)(*sort_cols)
        desc = (C.c_uint8 * k)(*[1 if x else 0 for x in (descending or [False] * k)])
        nf = (C.c_uint8 * k)(*[1 if x else 0 for x in (nulls_first or [False] * k)])
        out = C.c_void_p()
        self.ctx.check(lib().bgpu_batch_sort(self.ctx._h, self._h, cols, desc, nf, k, C.byref(out)))
        return DeviceBatch(self.ctx, out)

    def close(self):
        if getattr(self, "_h", None):
            lib().bgpu_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()


def create_hashes(ctx, dbatch, key_cols, out_device_ptr=None):
    """create_hashes over key columns; returns a numpy uint64 array unless out_device_ptr is given."""
    import torch  # plumbing: device scratch for the result

    n = dbatch.num_rows
    keys = (C.c_int32 * len(key_cols))(*key_cols)
    if out_device_ptr is not None:
        ctx.check(lib().bgpu_create_hashes(ctx._h, dbatch._h, keys, len(key_cols), C.c_void_p(out_device_ptr)))
        return None
    out = torch.empty(max(n, 1), dtype=torch.int64, device="cuda")
    ctx.check(lib().bgpu_create_hashes(ctx._h, dbatch._h, keys, len(key_cols), C.c_void_p(out.data_ptr())))
    ctx.synchronize()
    return out[:n].cpu().numpy().view(np.uint64)

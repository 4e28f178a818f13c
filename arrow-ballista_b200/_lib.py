"""Loader + ctypes signatures for libbgpu.so (the C ABI declared in include/bgpu.h).

There is no fallback: if the shared library is missing this raises, and on a box without a B200 every compute
call returns BGPU_ENODEVICE which `check()` turns into an exception.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BGPU_LIB") or os.path.join(_HERE, "libbgpu.so")  # BGPU_LIB: tuning builds only

BGPU_OK, BGPU_EINVAL, BGPU_EUNSUPPORTED, BGPU_ECUDA, BGPU_ENODEVICE, BGPU_EOVERFLOW, BGPU_ENCCL = range(7)
_ERR_NAMES = {1: "EINVAL", 2: "EUNSUPPORTED", 3: "ECUDA", 4: "ENODEVICE", 5: "EOVERFLOW", 6: "ENCCL"}


class BgpuError(RuntimeError):
    """Mirrors DataFusionError at the boundary (SURVEY.md §8b Errors): carries the category code."""

    def __init__(self, code, msg):
        super().__init__(f"libbgpu {_ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


class ArrowSchema(C.Structure):
    pass


class ArrowArray(C.Structure):
    pass


class ExprNode(C.Structure):
    """bgpu_expr_node (include/bgpu.h)."""
    _fields_ = [("op", C.c_int32), ("a", C.c_int32), ("precision", C.c_int32), ("scale", C.c_int32),
                ("lo", C.c_uint64), ("hi", C.c_uint64)]


ArrowSchema._fields_ = [
    ("format", C.c_char_p), ("name", C.c_char_p), ("metadata", C.c_char_p), ("flags", C.c_int64),
    ("n_children", C.c_int64), ("children", C.POINTER(C.POINTER(ArrowSchema))),
    ("dictionary", C.POINTER(ArrowSchema)), ("release", C.c_void_p), ("private_data", C.c_void_p)]
ArrowArray._fields_ = [
    ("length", C.c_int64), ("null_count", C.c_int64), ("offset", C.c_int64), ("n_buffers", C.c_int64),
    ("n_children", C.c_int64), ("buffers", C.POINTER(C.c_void_p)), ("children", C.POINTER(C.POINTER(ArrowArray))),
    ("dictionary", C.POINTER(ArrowArray)), ("release", C.c_void_p), ("private_data", C.c_void_p)]

# name -> (restype, argtypes); every symbol include/bgpu.h declares
SIGNATURES = {
    "bgpu_version": (C.c_char_p, []),
    "bgpu_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "bgpu_ctx_create_on_stream": (C.c_int, [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "bgpu_ctx_destroy": (None, [C.c_void_p]),
    "bgpu_last_error": (C.c_char_p, [C.c_void_p]),
    "bgpu_ctx_synchronize": (C.c_int, [C.c_void_p]),
    "bgpu_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "bgpu_ctx_kernel_launches": (C.c_uint64, [C.c_void_p]),
    "bgpu_ctx_enable_times": (C.c_int, [C.c_void_p, C.c_int32]),
    "bgpu_ctx_times": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.c_int32]),
    "bgpu_batch_upload": (C.c_int, [C.c_void_p, C.POINTER(ArrowArray), C.POINTER(ArrowSchema), C.POINTER(C.c_void_p)]),
    "bgpu_batch_wrap_device": (C.c_int, [C.c_void_p, C.POINTER(ArrowArray), C.POINTER(ArrowSchema),
                                         C.POINTER(C.c_void_p)]),
    "bgpu_batch_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(ArrowArray),
                                      C.POINTER(ArrowSchema)]),
    "bgpu_batch_export_device": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(ArrowArray), C.POINTER(ArrowSchema)]),
    "bgpu_batch_num_rows": (C.c_int64, [C.c_void_p]),
    "bgpu_batch_memory_size": (C.c_int64, [C.c_void_p]),
    "bgpu_batch_num_columns": (C.c_int32, [C.c_void_p]),
    "bgpu_batch_destroy": (None, [C.c_void_p]),
    "bgpu_create_hashes": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_void_p]),
    "bgpu_hash_partition": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_uint32,
                                      C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_void_p]),
    "bgpu_partition_offsets_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint32, C.c_void_p]),
    "bgpu_hash_partition_host": (C.c_int, [C.c_void_p, C.POINTER(ArrowArray), C.POINTER(ArrowSchema),
                                           C.POINTER(C.c_int32), C.c_int32, C.c_uint32, C.POINTER(ArrowArray),
                                           C.POINTER(ArrowSchema), C.POINTER(C.c_uint64)]),
    "bgpu_agg_create": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(ArrowSchema), C.POINTER(C.c_int32), C.c_int32,
                                  C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_char_p), C.c_int32,
                                  C.POINTER(C.c_void_p)]),
    "bgpu_agg_create_expr": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(ArrowSchema), C.POINTER(C.c_int32), C.c_int32,
                                       C.POINTER(C.c_int32), C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_char_p),
                                       C.c_int32, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]),
    "bgpu_agg_debug_compile": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(ArrowSchema), C.POINTER(C.c_int32), C.c_int32,
                                         C.POINTER(C.c_int32), C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_void_p,
                                         C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_char_p]),
    "bgpu_agg_last_path": (C.c_char_p, [C.c_void_p]),
    "bgpu_agg_hot_kernel_ms": (C.c_double, [C.c_void_p]),
    "bgpu_agg_update": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bgpu_agg_push_host": (C.c_int, [C.c_void_p, C.POINTER(ArrowArray), C.POINTER(ArrowSchema)]),
    "bgpu_agg_flush": (C.c_int, [C.c_void_p]),
    "bgpu_agg_ingest_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "bgpu_debug_pack_checked_host": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_uint32, C.c_void_p]),
    "bgpu_debug_pack_fixed_host": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int64, C.c_int32, C.POINTER(C.c_int64),
                                             C.POINTER(C.c_uint32), C.c_void_p]),
    "bgpu_agg_emit": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "bgpu_jit_stats": (None, [C.POINTER(C.c_uint64)]),
    "bgpu_agg_reset": (C.c_int, [C.c_void_p]),
    "bgpu_agg_num_groups": (C.c_int64, [C.c_void_p]),
    "bgpu_agg_destroy": (None, [C.c_void_p]),
    "bgpu_nccl_unique_id": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bgpu_nccl_comm_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "bgpu_nccl_comm_destroy": (None, [C.c_void_p]),
    "bgpu_exchange_plan": (C.c_uint64, [C.POINTER(C.c_uint64), C.c_int32, C.c_int32, C.c_uint32, C.POINTER(C.c_uint64),
                                        C.POINTER(C.c_uint32)]),
    "bgpu_all_to_all": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_uint64),
                                  C.c_uint32, C.POINTER(C.c_void_p)]),
    "bgpu_small_exchange": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_int32), C.c_int32,
                                      C.c_uint32, C.POINTER(C.c_void_p)]),
    "bgpu_peer_exchange_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_uint64, C.POINTER(C.c_void_p)]),
    "bgpu_partition_exchange": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_uint32,
                                          C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "bgpu_peer_exchange_window_bytes": (C.c_uint64, [C.c_void_p]),
    "bgpu_peer_exchange_destroy": (None, [C.c_void_p]),
    "bgpu_ipc_schema_message": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "bgpu_ipc_record_batch_metadata": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_int32, C.POINTER(C.c_int64),
                                                 C.c_int32, C.c_int64, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "bgpu_ipc_encode_partitions": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.c_uint32, C.c_int32,
                                             C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "bgpu_ipc_free": (None, [C.c_void_p, C.c_int]),
    "bgpu_ipc_reader_open": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]),
    "bgpu_ipc_reader_schema": (C.c_int, [C.c_void_p, C.POINTER(ArrowSchema)]),
    "bgpu_ipc_reader_next": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "bgpu_ipc_reader_close": (None, [C.c_void_p]),
    "bgpu_debug_lz4_frame_decode_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]),
    "bgpu_batch_sort": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.POINTER(C.c_uint8), C.c_int32,
                                  C.POINTER(C.c_void_p)]),
    "bgpu_batch_concat": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int32, C.POINTER(C.c_void_p)]),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                              f"g.build()'` (make -C arrow-ballista_b200/csrc). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError = header/library mismatch: fail loudly
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib

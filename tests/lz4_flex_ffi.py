"""TEST INFRASTRUCTURE: the reference's own LZ4 implementation, executed.

The reference compresses shuffle files with arrow-ipc's LZ4_FRAME codec, which is the `lz4_flex` crate
(/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:365-376 -> arrow-ipc 49 -> lz4_flex). No Rust
toolchain exists here, but this image ships a Rust cdylib that statically links lz4_flex-0.11.5 and keeps its local
symbols: wandb/bin/librust_parquet_ffi.so. Three of parquet-rs's codec methods in it are thin wrappers over lz4_flex and
are called here by address through ctypes (argument registers read off the disassembly; the first bytes of every function
are checked before it is called):

  LZ4Codec::decompress     = lz4_flex::frame::FrameDecoder::new(input).read(..) until EOF   (the reader arrow-ipc uses)
  LZ4RawCodec::compress    = lz4_flex::block::compress_into                                (the block compressor)
  LZ4RawCodec::decompress  = lz4_flex::block::decompress_into

`available()` is False on a machine without that exact library: tests that use this module skip there.
"""
import ctypes as C
import importlib.util
import os

_FUNCS = {  # name: (offset, first bytes)
    "create_reader": (0x370B40, None),
    "raw_compress": (0x66B6D0, "554889e54157415641554154534883e4c04881ec00010000"),
    "frame_decompress": (0x5FC520, "554157415641554154534881ec0010000048c70424000000"),
    "raw_decompress": (0x66B440, "554157415641554154534883ec384889fb4983f9010f85cd"),
}
_state = {}


class _Vec(C.Structure):  # alloc::vec::Vec<u8> as this rustc lays it out: {cap, ptr, len}
    _fields_ = [("cap", C.c_size_t), ("ptr", C.c_void_p), ("len", C.c_size_t)]


def _load():
    if "ok" in _state:
        return _state["ok"]
    _state["ok"] = False
    spec = importlib.util.find_spec("wandb")
    if spec is None or not spec.origin:
        return False
    path = os.path.join(os.path.dirname(spec.origin), "bin", "librust_parquet_ffi.so")
    if not os.path.exists(path):
        return False
    try:
        lib = C.CDLL(path)
        base = C.cast(lib.create_reader, C.c_void_p).value - _FUNCS["create_reader"][0]
    except (OSError, AttributeError):
        return False
    for name, (off, sig) in _FUNCS.items():
        if sig and C.string_at(base + off, len(sig) // 2).hex() != sig:
            return False  # another build of the library
    sret, vec = C.c_void_p, C.POINTER(_Vec)
    # (sret, &mut self, input.ptr, input.len, &mut Vec<u8>[, Option<usize> as (tag, value)])
    _state["raw_compress"] = C.CFUNCTYPE(C.c_void_p, sret, C.c_void_p, C.c_char_p, C.c_size_t, vec)(base + _FUNCS["raw_compress"][0])
    _state["frame_decompress"] = C.CFUNCTYPE(C.c_void_p, sret, C.c_void_p, C.c_char_p, C.c_size_t, vec, C.c_size_t, C.c_size_t)(
        base + _FUNCS["frame_decompress"][0])
    _state["raw_decompress"] = C.CFUNCTYPE(C.c_void_p, sret, C.c_void_p, C.c_char_p, C.c_size_t, vec, C.c_size_t, C.c_size_t)(
        base + _FUNCS["raw_decompress"][0])
    _state["lib"] = lib
    _state["ok"] = True
    return True


def available():
    return _load()


class Lz4FlexError(Exception):
    pass


def _call(fn, data, *extra):
    assert _load(), "lz4_flex FFI unavailable"
    result = (C.c_uint64 * 16)()   # Result<_, ParquetError>: tag 6 = Ok, payload in word 1
    out = _Vec(0, 1, 0)            # Vec::new(): no allocation, dangling pointer; grown by the library's own allocator
    this = C.c_uint64(0)           # the codecs are zero-sized or hold only a level
    fn(C.addressof(result), C.addressof(this), bytes(data), len(data), C.byref(out), *extra)
    if result[0] != 6:
        raise Lz4FlexError(f"lz4_flex returned an error (ParquetError tag {result[0]})")
    got = C.string_at(out.ptr, out.len) if out.len else b""
    return got, result[1]          # the Vec's buffer is left to leak: it belongs to the library's allocator


def frame_decompress(frame):
    """lz4_flex::frame::FrameDecoder over `frame`, read to the end."""
    got, total = _call(_state["frame_decompress"] if _load() else None, frame, 0, 0)
    assert total == len(got)
    return got


def block_compress(data):
    """lz4_flex::block::compress_into: one raw LZ4 block (no size prefix)."""
    return _call(_state["raw_compress"] if _load() else None, data)[0]


def block_decompress(block, size):
    """lz4_flex::block::decompress_into a buffer of exactly `size` bytes."""
    return _call(_state["raw_decompress"] if _load() else None, block, 1, size)[0]

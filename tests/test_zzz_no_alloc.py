"""bgpu_partition_offsets_noalloc (the allocation-free variant of the partitioner's histogram + scan; static __device__
scratch, n_partitions <= 512) through the C ABI against bgpu_hash_partition's offsets, bit-exact, on two calls in a row.

Collected last on purpose (a failure here must not hide another suite)."""
import ctypes as C

import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu


def _reference_offsets(bb, ctx, keys, n_partitions):
    """out_offsets of the existing entry point (bgpu_hash_partition) over the same key column."""
    schema = pa.schema([("k", pa.int64())])
    d = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=keys.data_ptr())], keys.numel(), keepalive=keys)
    _, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], n_partitions), ctx).partition_device(d)
    return [int(x) for x in offs]


def _noalloc_offsets(bb, ctx, keys, n_partitions, out):
    from ballista_b200._lib import lib

    ctx.check(lib().bgpu_partition_offsets_noalloc(ctx._h, C.c_void_p(keys.data_ptr()), keys.numel(), n_partitions,
                                                   C.c_void_p(out.data_ptr())))
    ctx.synchronize()
    return out[: n_partitions + 1].cpu().tolist()


def test_noalloc_offsets_match_hash_partition_two_calls_in_a_row(bb, ctx):
    import torch

    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(5)
    # caller-owned input and output buffers, allocated before the calls under test
    k1 = torch.randint(-2**62, 2**62, (3_000_017,), device=dev, generator=g)
    k2 = torch.randint(0, 1000, (70_001,), device=dev, generator=g)  # fewer rows, fewer chunks, fewer partitions: the
    out = torch.full((513,), -1, dtype=torch.int64, device=dev)     # second call must not see the first call's scratch
    want1, want2 = _reference_offsets(bb, ctx, k1, 512), _reference_offsets(bb, ctx, k2, 7)
    # the first use of the entry point loads its CUDA module (lazy loading): that is when the device materialises the
    # static scratch array, not an allocation by the call
    _noalloc_offsets(bb, ctx, k2, 3, out)
    ctx.synchronize()
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(dev)  # device-wide: sees this library's allocations too
    got1 = _noalloc_offsets(bb, ctx, k1, 512, out)
    got2 = _noalloc_offsets(bb, ctx, k2, 7, out)
    assert got1 == want1
    assert got2 == want2
    assert got1[-1] == k1.numel() and got2[-1] == k2.numel()
    assert torch.cuda.mem_get_info(dev)[0] >= free_before  # the two calls took no device memory
    # the same call again (the scratch keeps the last call's contents) and an empty input
    assert _noalloc_offsets(bb, ctx, k1, 512, out) == want1
    assert _noalloc_offsets(bb, ctx, k1[:0], 16, out) == [0] * 17


def test_noalloc_rejects_sizes_beyond_the_static_scratch(bb, ctx):
    import torch

    from ballista_b200._lib import BGPU_EINVAL, lib

    k = torch.zeros(16, dtype=torch.int64, device="cuda:0")
    out = torch.zeros(1024, dtype=torch.int64, device="cuda:0")
    rc = lib().bgpu_partition_offsets_noalloc(ctx._h, C.c_void_p(k.data_ptr()), 16, 513, C.c_void_p(out.data_ptr()))
    assert rc == BGPU_EINVAL
    assert b"maximum of 512" in lib().bgpu_last_error(ctx._h)

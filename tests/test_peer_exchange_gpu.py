"""GPU parity of the fused hash repartition + exchange (bgpu_partition_exchange: the scatter kernel stores runs straight
into the owner ranks' receive windows over NVLink peer memory). The received batch must be, bit for bit, what
bgpu_all_to_all(bgpu_hash_partition(batch)) gives and what the oracle's BatchPartitioner says: this rank's partitions of
every rank, in (partition, source rank) order, input order kept inside.

world = 1 runs on any box (the window is local: it exercises the kernel, the layout and the window growth); world = 2 / 4
need that many GPUs (`gpurun --gpus N`) and are skipped otherwise."""
import os
import sys

import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_PARTITIONS = 16
KEYS = [0]


def make_input(rank, n, nulls=False, seed=0):
    """An aggregate-state shaped batch: Int64 key, two Decimal128 sums, a count, plus 4-byte and float columns."""
    import decimal

    rng = np.random.default_rng(100 * seed + 7 + rank)
    dec = [decimal.Decimal(int(v)) / 100 for v in rng.integers(-10**14, 10**14, n)]
    return pa.RecordBatch.from_arrays(
        [pa.array(rng.integers(0, 1 << 40, n), type=pa.int64()),
         pa.array(dec, type=pa.decimal128(25, 2)),
         pa.array(dec[::-1], type=pa.decimal128(25, 2)),
         pa.array(rng.integers(1, 8, n), type=pa.int64(), mask=(rng.random(n) < 0.1) if nulls else None),
         pa.array(rng.integers(-2**31, 2**31 - 1, n).astype("int32"), type=pa.int32()),
         pa.array(rng.random(n))],
        names=["k", "s1", "s2", "c", "i", "f"])


def expected_received(oracle, rank, world, sizes, nulls, seed):
    per_source = []
    for s in range(world):
        parts, _ = oracle.partition(make_input(s, sizes[s], nulls, seed), KEYS, N_PARTITIONS)
        per_source.append({p: b for p, b in parts})
    chunks = [per_source[s][p] for p in range(rank, N_PARTITIONS, world) for s in range(world) if p in per_source[s]]
    schema = make_input(0, 1, nulls, seed).schema
    return pa.Table.from_batches(chunks, schema=schema).combine_chunks() if chunks else schema.empty_table()


def run_rank(bb, oracle, ctx, rank, world):
    ex = bb.Exchange(ctx, rank, world)
    # rounds: small, a larger one that outgrows the 1 MB window (collective regrowth), an unbalanced one with an empty
    # rank, then the same window reused by a small batch again; last: nulls -> every rank falls back to NCCL together
    rounds = [([3000] * world, False), ([150_000 + 1000 * r for r in range(world)], False),
              ([0 if r == world - 1 else 20_000 for r in range(world)], False), ([777] * world, False), ([5000] * world, True)]
    for seed, (sizes, nulls) in enumerate(rounds):
        dev = bb.DeviceBatch.upload(ctx, make_input(rank, sizes[rank], nulls, seed))
        got = ex.partition_exchange(dev, KEYS, N_PARTITIONS, window_bytes=1 << 20).download()
        assert ex.last_path == ("nccl" if nulls else "fused_peer"), ex.last_path
        want = expected_received(oracle, rank, world, sizes, nulls, seed)
        assert got.num_rows == want.num_rows, (seed, got.num_rows, want.num_rows)
        assert pa.Table.from_batches([got], schema=want.schema).equals(want), f"rank {rank} round {seed}: differs from the oracle"
        if not nulls:
            # and from the two-step path it replaces
            part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash(KEYS, N_PARTITIONS), ctx).partition_device(dev)
            two_step = ex.all_to_all(part, offs, N_PARTITIONS).download() if world > 1 else part.download()
            assert two_step.equals(got)
            sent = sum(b.num_rows for p, b in oracle.partition(make_input(rank, sizes[rank], nulls, seed), KEYS, N_PARTITIONS)[0]
                       if p % world != rank) * (8 + 16 + 16 + 8 + 4 + 8)
            assert ex.last_sent_bytes == sent
    # Leases: the result batch of a fused exchange IS its receive window (no copy-out), so it holds the window until it
    # dies; the next exchange lands in the second window; with both windows held the call answers "unsupported" on
    # every rank together (the two-step path delivers the same rows); a released window is used again.
    def exchange(sizes, seed):
        dev = bb.DeviceBatch.upload(ctx, make_input(rank, sizes[rank], False, seed))
        return ex.partition_exchange(dev, KEYS, N_PARTITIONS, window_bytes=1 << 20), expected_received(oracle, rank, world, sizes, False, seed)

    def same(batch, want):
        return pa.Table.from_batches([batch.download()], schema=want.schema).equals(want)

    held_a, want_a = exchange([4000 + 10 * r for r in range(world)], 11)
    assert ex.last_path == "fused_peer" and same(held_a, want_a)
    held_b, want_b = exchange([90_000] * world, 12)            # the second window (made and grown on demand)
    assert ex.last_path == "fused_peer" and same(held_b, want_b)
    assert same(held_a, want_a), "the first result was overwritten while it was still held"
    third, want_c = exchange([2500] * world, 13)               # both windows are leased: everybody takes the two-step path
    assert ex.last_path == "nccl" and same(third, want_c)
    assert same(held_a, want_a) and same(held_b, want_b)
    del held_a
    fourth, want_d = exchange([3100] * world, 14)              # the released window serves again
    assert ex.last_path == "fused_peer" and same(fourth, want_d) and same(held_b, want_b)
    del held_b, third, fourth
    ex.close()


def test_partition_exchange_single_rank(bb, oracle, ctx):
    run_rank(bb, oracle, ctx, 0, 1)


def _worker(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist

    import ballista_b200 as bb
    import oracle as O

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # rendezvous only: carries the ncclUniqueId
    try:
        ctx = bb.Context(rank)
        run_rank(bb, O, ctx, rank, world)
        ctx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_partition_exchange_multi_gpu(world, oracle):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mp.spawn(_worker, args=(world, 31700 + os.getpid() % 2000), nprocs=world, join=True)

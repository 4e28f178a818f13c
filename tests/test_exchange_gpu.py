"""GPU parity of bgpu_all_to_all (NCCL over NVLink): the received batch must hold exactly this rank's partitions of
every rank's BatchPartitioner output, in (partition, source rank) order, bit for bit (values, validity, utf8).

world = 1 runs on any box (NCCL send/recv to self: exercises the header all-gather, validity byte staging and utf8
offset rebuild); world = 2 needs two GPUs (`gpurun --gpus 2`) and is skipped otherwise.
"""
import os
import sys

import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_PARTITIONS = 6
KEYS = [0, 1]


def make_input(rank, n=20000):
    rng = np.random.default_rng(7 + rank)
    words = ["", "a", "bc", "def", "ghij", "klmnopqr", "stuvwxyz0"]
    return pa.RecordBatch.from_arrays(
        [pa.array(rng.integers(0, 500, n), type=pa.int64()),
         pa.array([words[i] for i in rng.integers(0, len(words), n)], type=pa.string(), mask=rng.random(n) < 0.05),
         pa.array(rng.integers(-10**9, 10**9, n), type=pa.int64(), mask=rng.random(n) < 0.1),
         pa.array(rng.random(n)),
         pa.array(rng.integers(-2**31, 2**31 - 1, n).astype("int32"), type=pa.int32())],
        names=["k", "s", "v", "f", "i"])


def expected_received(oracle, rank, world):
    """Concatenation over owned partitions p (ascending) and sources s (ascending) of the oracle's partition p of s."""
    per_source = []
    for s in range(world):
        parts, _ = oracle.partition(make_input(s), KEYS, N_PARTITIONS)
        per_source.append({p: b for p, b in parts})
    chunks = []
    for p in range(rank, N_PARTITIONS, world):
        for s in range(world):
            if p in per_source[s]:
                chunks.append(per_source[s][p])
    schema = make_input(0).schema
    return pa.Table.from_batches(chunks, schema=schema) if chunks else schema.empty_table()


def run_rank(bb, oracle, ctx, rank, world):
    ex = bb.Exchange(ctx, rank, world)
    part = bb.BatchPartitioner.try_new(bb.Partitioning.Hash(KEYS, N_PARTITIONS), ctx)
    dev, offsets = part.partition_device(bb.DeviceBatch.upload(ctx, make_input(rank)))
    got = ex.all_to_all(dev, offsets, N_PARTITIONS).download()
    want = expected_received(oracle, rank, world)
    assert got.num_rows == want.num_rows
    assert pa.Table.from_batches([got]).equals(want.combine_chunks()), f"rank {rank}: received batch differs from the oracle"
    ex.close()


def make_small_input(rank, n):
    """An aggregate-state-like batch of a few rows: utf8 + Int64 keys (nulls on some ranks only), Decimal128 / Int64 / Float64
    payloads, a bool."""
    import decimal

    rng = np.random.default_rng(100 + rank)
    flags = ["A", "N", "R", "", "OF"]
    mask = rng.random(n) < 0.2 if rank % 2 else None
    return pa.RecordBatch.from_arrays(
        [pa.array([flags[i] for i in rng.integers(0, len(flags), n)], type=pa.string(), mask=mask),
         pa.array(rng.integers(0, 9, n), type=pa.int64()),
         pa.array([decimal.Decimal(int(x)).scaleb(-2) for x in rng.integers(-10**15, 10**15, n)], type=pa.decimal128(25, 2)),
         pa.array(rng.integers(0, 10**9, n), type=pa.int64(), mask=rng.random(n) < 0.3),
         pa.array(rng.random(n)), pa.array(rng.random(n) < 0.5)],
        names=["rf", "k", "sum", "cnt", "f", "b"])


def run_small_rank(bb, oracle, ctx, rank, world):
    """bgpu_small_exchange == the rows all_to_all(hash_partition(batch)) delivers (as a multiset: the order is by source)."""
    ex = bb.Exchange(ctx, rank, world)
    for n_parts, rows in ((16, 5), (6, 200), (7, 0)):
        inputs = [make_small_input(s, rows if s != 1 else max(0, rows - 3)) for s in range(world)]
        got = ex.small_exchange(bb.DeviceBatch.upload(ctx, inputs[rank]), [0, 1], n_parts)
        assert got is not None and ex.last_path == "small_allgather"
        chunks = []
        for s in range(world):
            parts, _ = oracle.partition(inputs[s], [0, 1], n_parts)
            chunks += [b for p, b in parts if p % world == rank]
        want = pa.Table.from_batches(chunks, schema=inputs[0].schema) if chunks else inputs[0].schema.empty_table()
        gt = pa.Table.from_batches([got.download()])
        assert gt.num_rows == want.num_rows, (rank, n_parts, rows)
        keys = [(nm, "ascending") for nm in ("k", "sum", "f")]
        assert gt.sort_by(keys).equals(want.combine_chunks().sort_by(keys)), f"rank {rank}: small exchange differs ({n_parts}, {rows})"
    # a batch beyond the small path on ONE rank: every rank is told to take the general path
    big = make_small_input(rank, 300 if rank == world - 1 else 4)
    assert ex.small_exchange(bb.DeviceBatch.upload(ctx, big), [0, 1], 16) is None
    ex.close()


def test_small_exchange_single_rank(bb, oracle, ctx):
    run_small_rank(bb, oracle, ctx, 0, 1)


def test_all_to_all_single_rank(bb, oracle, ctx):
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
        run_small_rank(bb, O, ctx, rank, world)
        ctx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_all_to_all_multi_gpu(world, oracle):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mp.spawn(_worker, args=(world, 29700 + os.getpid() % 2000), nprocs=world, join=True)

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

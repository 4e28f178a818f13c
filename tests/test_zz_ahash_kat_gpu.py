"""The device partitioner against hashes produced by EXECUTED upstream ahash-0.8.12 (tests/golden/ahash_0_8_12_kat.json,
oracle/evidence/make_ahash_kat.py): over a Binary key column every row lands in partition `executed hash % n`, in input
order inside a partition (BatchPartitioner's hash arm,
/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-391), for n = 2, 16 and 512."""
import json
import os

import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ahash_0_8_12_kat.json")


@pytest.mark.parametrize("n", [2, 16, 512])
def test_gpu_partitioning_follows_the_executed_hashes(bb, ctx, n):
    vs = [v for v in json.load(open(GOLDEN))["hash_one_bytes"] if v["keys"][0] == "452821e638d01377"]  # seeds (0,0,0,0)
    keys = [bytes.fromhex(v["data"]) for v in vs]
    hashes = [int(v["hash"], 16) for v in vs]
    batch = pa.RecordBatch.from_arrays([pa.array(keys, type=pa.binary()), pa.array(range(len(keys)), type=pa.int64())],
                                       names=["k", "row"])
    part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], n), ctx).partition_device(bb.DeviceBatch.upload(ctx, batch))
    got = part.download()
    assert offs[0] == 0 and offs[n] == len(keys) and got.num_rows == len(keys)
    rows = got.column("row").to_pylist()
    for p in range(n):
        seg = rows[offs[p]:offs[p + 1]]
        assert seg == sorted(seg) and all(hashes[r] % n == p for r in seg), p
    assert got.column("k").to_pylist() == [keys[r] for r in rows]


def test_gpu_create_hashes_equal_the_executed_hasher(bb, ctx):
    """create_hashes over a Binary column (DataFusion hashes `&[u8]` values with hash_one) = what the compiled upstream
    hasher gave for the same bytes with the keys seeds (0,0,0,0) store; a leading null keeps the zero the buffer starts with."""
    vs = [v for v in json.load(open(GOLDEN))["hash_one_bytes"] if v["keys"][0] == "452821e638d01377"]
    keys = [bytes.fromhex(v["data"]) for v in vs]
    want = [int(v["hash"], 16) for v in vs]
    batch = pa.RecordBatch.from_arrays([pa.array(keys, type=pa.binary())], names=["k"])
    assert bb.create_hashes(ctx, bb.DeviceBatch.upload(ctx, batch), [0]).tolist() == want
    with_null = pa.RecordBatch.from_arrays([pa.array([None] + keys, type=pa.binary())], names=["k"])
    assert bb.create_hashes(ctx, bb.DeviceBatch.upload(ctx, with_null), [0]).tolist() == [0] + want

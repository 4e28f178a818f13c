"""N > 1 host logic of the exchange on CPU: two `gloo` ranks partition their own input with the ORACLE (the checker),
agree on the received layout through libbgpu's bgpu_exchange_plan (pure host code of the product), move the rows with
torch.distributed, and the per-rank FinalPartitioned results must add up to the single-process answer.

Reference behaviour restated: output partition p of every map task is read by the one reduce task that owns p
(/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:226-298); here the owner is rank p % world.
"""
import os
import sys

import numpy as np
import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_PARTITIONS = 5
WORLD = 2


def make_input(rank):
    rng = np.random.default_rng(100 + rank)
    n = 3000 + 500 * rank
    return pa.RecordBatch.from_arrays(
        [pa.array(rng.integers(0, 40, n), type=pa.int64()),
         pa.array([f"g{int(i)}" for i in rng.integers(0, 7, n)], type=pa.string(), mask=rng.random(n) < 0.05),
         pa.array(rng.integers(-1000, 1000, n), type=pa.int64(), mask=rng.random(n) < 0.1)], names=["k", "s", "v"])


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist

    import ballista_b200 as bb
    import oracle as O

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        batch = make_input(rank)
        parts, _ = O.partition(batch, [0, 1], N_PARTITIONS)  # [(p, RecordBatch)], empty partitions skipped
        by_p = {p: b for p, b in parts}
        offsets = [0]
        for p in range(N_PARTITIONS):
            offsets.append(offsets[-1] + (by_p[p].num_rows if p in by_p else 0))
        all_offsets = [None] * world
        dist.all_gather_object(all_offsets, offsets)
        recv_off, owned = bb.exchange_plan(all_offsets, rank, world, N_PARTITIONS)
        assert owned == [p for p in range(N_PARTITIONS) if bb.owner_of(p, world) == rank]
        # the "wire": every rank publishes its partitions, each rank picks what it owns in (partition, source) order
        published = [None] * world
        dist.all_gather_object(published, {p: b.to_pylist() for p, b in by_p.items()})
        rows = []
        for k, p in enumerate(owned):
            for s in range(world):
                assert recv_off[k * world + s] == len(rows)
                rows.extend(published[s].get(p, []))
        assert recv_off[-1] == len(rows)
        received = pa.RecordBatch.from_pylist(rows, schema=batch.schema)
        # every received row belongs to a partition this rank owns
        for p, b in O.partition(received, [0, 1], N_PARTITIONS)[0]:
            assert bb.owner_of(p, world) == rank
        with open(os.path.join(out_dir, f"recv{rank}.arrow"), "wb") as f:
            with pa.ipc.new_stream(f, batch.schema) as w:
                w.write_batch(received)
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_layout_and_union(tmp_path):
    import torch.multiprocessing as mp

    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(WORLD, port, str(tmp_path)), nprocs=WORLD, join=True)
    got = []
    for r in range(WORLD):
        with pa.ipc.open_stream(str(tmp_path / f"recv{r}.arrow")) as rd:
            got.extend(rd.read_all().to_pylist())
    want = []
    for r in range(WORLD):
        want.extend(make_input(r).to_pylist())
    key = lambda d: (d["k"], d["s"] or "", d["s"] is None, d["v"] if d["v"] is not None else -10**9)
    assert sorted(got, key=key) == sorted(want, key=key)  # nothing lost, nothing duplicated


def test_exchange_plan_edge_cases():
    sys.path.insert(0, ROOT)
    import ballista_b200 as bb

    # more ranks than partitions: the extra rank owns nothing
    recv, owned = bb.exchange_plan([[0, 3], [0, 0], [0, 2]], 1, 3, 1)
    assert owned == [] and recv == [0]
    recv, owned = bb.exchange_plan([[0, 3], [0, 0], [0, 2]], 0, 3, 1)
    assert owned == [0] and recv == [0, 3, 3, 5]
    # empty partitions keep equal offsets
    recv, owned = bb.exchange_plan([[0, 0, 0, 4], [0, 0, 1, 1]], 0, 2, 3)
    assert owned == [0, 2] and recv == [0, 0, 0, 4, 4]

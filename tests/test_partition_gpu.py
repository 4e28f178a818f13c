"""GPU parity: create_hashes and BatchPartitioner (libbgpu, through the C ABI) against the oracle, bit-exact."""
import numpy as np
import pyarrow as pa
import pytest

from helpers import golden_hash_cases, random_batch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,arrays,expected", golden_hash_cases(), ids=[c[0] for c in golden_hash_cases()])
def test_gpu_hashes_match_golden(bb, ctx, name, arrays, expected):
    b = pa.RecordBatch.from_arrays(arrays, names=[f"c{i}" for i in range(len(arrays))])
    d = bb.DeviceBatch.upload(ctx, b)
    got = bb.create_hashes(ctx, d, list(range(len(arrays))))
    assert got.tolist() == expected


@pytest.mark.parametrize("keys", [["k64"], ["k32"], ["s"], ["d"], ["dec"], ["b"], ["f"], ["u8", "i16"],
                                  ["flag", "small"], ["k64", "s", "d", "dec", "date"]])
def test_gpu_hashes_match_oracle_random(bb, ctx, oracle, keys):
    rng = np.random.default_rng(11)
    b = random_batch(rng, 10_000)
    idx = [b.schema.names.index(k) for k in keys]
    want = oracle.create_hashes([b.column(i) for i in idx])
    got = bb.create_hashes(ctx, bb.DeviceBatch.upload(ctx, b), idx)
    assert (got == want).all()


def _check_partition(bb, ctx, oracle, b, key_idx, n):
    got = []
    bb.BatchPartitioner.try_new(bb.Partitioning.Hash(key_idx, n), ctx).partition(b, lambda p, rb: got.append((p, rb)))
    want, _ = oracle.partition(b, key_idx, n)
    assert [p for p, _ in got] == [p for p, _ in want]
    for (p, g), (_, w) in zip(got, want):
        assert g.schema.equals(w.schema), (g.schema, w.schema)
        for name, gc, wc in zip(g.schema.names, g.columns, w.columns):
            assert gc.equals(wc), f"partition {p} column {name} differs"


@pytest.mark.parametrize("n", [1, 2, 3, 8, 16, 64, 200, 512, 2048])
def test_partition_matches_oracle(bb, ctx, oracle, n):
    b = random_batch(np.random.default_rng(n), 30_000)
    _check_partition(bb, ctx, oracle, b, [0], n)


@pytest.mark.parametrize("keys", [["s"], ["d", "small"], ["flag", "k32"], ["dec"]])
def test_partition_other_keys(bb, ctx, oracle, keys):
    b = random_batch(np.random.default_rng(3), 20_000)
    _check_partition(bb, ctx, oracle, b, [b.schema.names.index(k) for k in keys], 16)


@pytest.mark.parametrize("rows", [0, 1, 31, 32, 33, 4095, 4096, 4097, 8193])
def test_partition_ragged_sizes(bb, ctx, oracle, rows):
    b = random_batch(np.random.default_rng(rows), rows)
    _check_partition(bb, ctx, oracle, b, [0], 16)


def test_partition_sliced_input(bb, ctx, oracle):
    b = random_batch(np.random.default_rng(5), 5000).slice(13, 3001)
    _check_partition(bb, ctx, oracle, b, [0, 9], 7)


def test_partition_skew_single_partition(bb, ctx, oracle):
    n = 50_000
    b = pa.RecordBatch.from_arrays([pa.array(np.full(n, 42), type=pa.int64()), pa.array(np.arange(n))], names=["k", "v"])
    _check_partition(bb, ctx, oracle, b, [0], 64)


def test_partition_large_is_stable_and_complete(bb, ctx, oracle):
    """Size-independent properties at a size the oracle does not check row by row."""
    import torch
    n, parts = 20_000_000, 64
    g = torch.Generator(device="cuda").manual_seed(1)
    k = torch.randint(-2**62, 2**62, (n,), device="cuda", generator=g)
    v = torch.arange(n, device="cuda")
    schema = pa.schema([("k", pa.int64()), ("row", pa.int64())])
    d = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=k.data_ptr()), dict(values=v.data_ptr())], n, (k, v))
    out, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], parts), ctx).partition_device(d)
    assert offs[0] == 0 and offs[-1] == n and all(a <= b for a, b in zip(offs, offs[1:]))
    for p in (0, 17, parts - 1):
        rb = out.download(offs[p], offs[p + 1] - offs[p])
        rows = rb.column(1).to_numpy()
        assert (np.diff(rows) > 0).all()  # stable
        hs = oracle.create_hashes([rb.column(0)])
        assert ((hs % parts) == p).all()  # every row belongs here
        assert (rb.column(0).to_numpy() == k[torch.from_numpy(rows).cuda()].cpu().numpy()).all()  # payload follows key
    whole = out.download()
    assert np.array_equal(np.sort(whole.column(1).to_numpy()), np.arange(n))  # a permutation: nothing lost


def test_unsupported_scheme_is_an_error(bb, ctx):
    with pytest.raises(bb.BgpuError):
        bb.BatchPartitioner.try_new(bb.Partitioning.RoundRobinBatch(4), ctx)
    with pytest.raises(bb.BgpuError):
        bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 5000), ctx).partition(
            random_batch(np.random.default_rng(0), 10), lambda p, b: None)


def test_context_time_metrics_account_h2d_kernel_d2h(bb):
    """bgpu_ctx_enable_times / bgpu_ctx_times: the h2d_time / kernel_time / d2h_time a GPU operator reports next to the
    reference's write_time / repart_time (shuffle_writer.rs:139-163). Off by default; each class grows only with its own
    kind of call; reset zeroes the sums."""
    ctx = bb.Context(0)
    rng = np.random.default_rng(3)
    b = pa.RecordBatch.from_arrays([pa.array(rng.integers(0, 1 << 40, 200_000), type=pa.int64()),
                                    pa.array(rng.random(200_000))], names=["k", "f"])
    d = bb.DeviceBatch.upload(ctx, b)
    assert ctx.times() == {"h2d_time": 0.0, "kernel_time": 0.0, "d2h_time": 0.0}   # not enabled: nothing is recorded
    ctx.enable_times(True)
    d = bb.DeviceBatch.upload(ctx, b)
    t1 = ctx.times()
    assert t1["h2d_time"] > 0 and t1["kernel_time"] == 0 and t1["d2h_time"] == 0, t1
    part = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 8), ctx)
    out, offs = part.partition_device(d)
    t2 = ctx.times()
    assert t2["kernel_time"] > 0 and t2["h2d_time"] == t1["h2d_time"] and t2["d2h_time"] == 0, t2
    assert out.download().num_rows == b.num_rows
    t3 = ctx.times(reset=True)
    assert t3["d2h_time"] > 0 and t3["kernel_time"] == t2["kernel_time"], t3
    assert ctx.times() == {"h2d_time": 0.0, "kernel_time": 0.0, "d2h_time": 0.0}
    ctx.close()

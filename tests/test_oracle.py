"""CPU tests: the oracle against the golden vectors and its own invariants (no GPU needed)."""
import numpy as np
import pyarrow as pa
import pytest

from helpers import golden_hash_cases, random_batch


@pytest.mark.parametrize("name,arrays,expected", golden_hash_cases(), ids=[c[0] for c in golden_hash_cases()])
def test_oracle_hashes_match_golden(oracle, name, arrays, expected):
    got = oracle.create_hashes(arrays)
    assert got.tolist() == expected


def test_reference_ignored_test_inputs(oracle):
    # /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:611-668 (ignored): a=UInt32[1,2],
    # Hash(a, 2). With the ahash-0.8 fallback hasher both land in partition 1 (oracle/evidence/README.md).
    h = oracle.create_hashes([pa.array([1, 2], type=pa.uint32())])
    assert h.tolist() == [0xA21258D088C87A13, 0xC2CE0F0A357699BD]
    assert (h % 2).tolist() == [1, 1]


def test_null_rows_keep_running_hash(oracle):
    a = pa.array([1, None, 3], type=pa.int64())
    b = pa.array(["x", "y", None])
    h = oracle.create_hashes([a, b])
    hb = oracle.create_hashes([b])
    ha = oracle.create_hashes([a])
    assert h[1] == ((17 * 37 + int(hb[1])) * 37 + 0) % 2**64  # null first column leaves 0, then combine
    assert h[2] == ha[2]  # null second column leaves the running hash
    assert oracle.create_hashes([pa.array([None, None], type=pa.int64())]).tolist() == [0, 0]


def test_sliced_arrays_hash_like_copies(oracle):
    rng = np.random.default_rng(1)
    b = random_batch(rng, 500)
    for i, col in enumerate(b.columns):
        whole = oracle.create_hashes([col])
        part = oracle.create_hashes([col.slice(37, 200)])
        assert (whole[37:237] == part).all(), b.schema.names[i]


@pytest.mark.parametrize("n_part", [1, 2, 3, 16, 64, 512])
def test_partition_properties(oracle, n_part):
    rng = np.random.default_rng(n_part)
    b = random_batch(rng, 3000)
    parts, hashes = oracle.partition(b, [0, 9], n_part)
    assert sum(rb.num_rows for _, rb in parts) == b.num_rows
    assert [p for p, _ in parts] == sorted(p for p, _ in parts)
    tbl = pa.Table.from_batches([b]).append_column("__row", pa.array(np.arange(b.num_rows)))
    parts2, _ = oracle.partition(tbl.to_batches()[0], [0, 9], n_part)
    for p, rb in parts2:
        rows = rb.column(rb.num_columns - 1).to_numpy()
        assert (np.diff(rows) > 0).all()  # input order preserved inside a partition
        assert ((hashes[rows] % n_part) == p).all()
        assert rb.slice(0, rb.num_rows).drop_columns(["__row"]).equals(b.take(pa.array(rows)))


def test_empty_batch(oracle):
    b = random_batch(np.random.default_rng(0), 0)
    parts, hashes = oracle.partition(b, [0], 8)
    assert parts == [] and len(hashes) == 0


def test_grouped_aggregate_matches_pyarrow(oracle):
    rng = np.random.default_rng(5)
    b = random_batch(rng, 4000)
    names = b.schema.names
    gi = [names.index("flag"), names.index("small")]
    aggs = [("sum", names.index("dec"), "sd"), ("avg", names.index("f"), "af"), ("count", None, "c"),
            ("count", names.index("k32"), "ck"), ("sum", names.index("k32"), "sk"), ("avg", names.index("dec"), "ad")]
    P = [oracle.GroupedAggregate("partial", gi, aggs) for _ in range(3)]
    for i in range(0, b.num_rows, 512):
        P[(i // 512) % 3].update(b.slice(i, 512))
    F = oracle.GroupedAggregate("final", [0, 1], aggs,
                                arg_types=[b.column(a).type if a is not None else None for _, a, _ in aggs])
    for p in P:
        F.update(p.emit())
    res = pa.Table.from_batches([F.emit()]).sort_by([("flag", "ascending"), ("small", "ascending")])
    want = pa.Table.from_batches([b]).group_by(["flag", "small"]).aggregate(
        [("dec", "sum"), ("f", "mean"), ([], "count_all"), ("k32", "count"), ("k32", "sum")]).sort_by(
        [("flag", "ascending"), ("small", "ascending")])
    assert res["sd"].cast(pa.decimal128(38, 2)).equals(want["dec_sum"].cast(pa.decimal128(38, 2)))
    assert np.allclose(res["af"].to_numpy(), want["f_mean"].to_numpy(), rtol=1e-12)
    assert res["c"].to_pylist() == want["count_all"].to_pylist()
    assert res["ck"].to_pylist() == want["k32_count"].to_pylist()
    assert res["sk"].to_pylist() == want["k32_sum"].to_pylist()
    # decimal avg: truncating (sum * 10^4) / count at scale s+4
    sums = want["dec_sum"].to_pylist()
    cnts = pa.Table.from_batches([b]).group_by(["flag", "small"]).aggregate([("dec", "count")]).sort_by(
        [("flag", "ascending"), ("small", "ascending")])["dec_count"].to_pylist()
    for got, s, c in zip(res["ad"].to_pylist(), sums, cnts):
        unscaled = int(s.scaleb(2)) * 10**4
        q = abs(unscaled) // c * (1 if unscaled >= 0 else -1)
        assert int(got.scaleb(6)) == q

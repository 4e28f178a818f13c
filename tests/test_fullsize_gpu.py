"""Size-independent properties at sizes the oracle does not check row by row (the prompt's "full size" leg):
the GPU hot path against independent exact integer arithmetic in torch on the same device-resident input."""
import decimal
import os
import sys

import numpy as np
import pyarrow as pa
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

pytestmark = pytest.mark.gpu


def test_q1_sf10_shard_against_torch_integer_sums(bb, ctx):
    """TPC-H Q1 over 60 M synthetic lineitem rows (SF10) through Partial -> repartition -> Final: every Decimal128 sum,
    the counts and the truncating decimal averages equal exact int64 arithmetic done by torch (all Q1 sums at this size
    fit int64: max charge 1.05e7 * 100 * 108 * 6e7 < 2^63)."""
    import torch

    import bench

    n = 60_000_000
    cols = bench.gen_lineitem_torch(n, 42, "cuda")
    batch = bench.wrap_lineitem(bb, ctx, cols, n)
    hp = bench.HotPath(bb, ctx)
    out = pa.Table.from_batches([hp.step([batch]).download()])
    assert hp.last_path == "fused_private"
    qty, price, disc, tax = (cols[k][:, 0] for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax"))
    keep = cols["l_shipdate"] <= bench.Q1_DATE
    disc_price = price * (100 - disc)
    charge = disc_price * (100 + tax)
    got = {(r["l_returnflag"], r["l_linestatus"]): r for r in out.to_pylist()}
    seen_rows = 0
    for rf in "ARN":
        for ls in "FO":
            m = keep & (cols["rf_data"] == ord(rf)) & (cols["ls_data"] == ord(ls))
            cnt = int(m.sum())
            if cnt == 0:
                assert (rf, ls) not in got
                continue
            seen_rows += cnt
            r = got[(rf, ls)]
            sums = {k: int(v[m].sum()) for k, v in (("sum_qty", qty), ("sum_base_price", price),
                                                    ("sum_disc_price", disc_price), ("sum_charge", charge))}
            assert r["count_order"] == cnt
            assert r["sum_qty"] == decimal.Decimal(sums["sum_qty"]).scaleb(-2)
            assert r["sum_base_price"] == decimal.Decimal(sums["sum_base_price"]).scaleb(-2)
            assert r["sum_disc_price"] == decimal.Decimal(sums["sum_disc_price"]).scaleb(-4)
            assert r["sum_charge"] == decimal.Decimal(sums["sum_charge"]).scaleb(-6)
            # Decimal128 avg = (sum * 10^4) / count, truncating (scale + 4)
            assert r["avg_qty"] == decimal.Decimal(sums["sum_qty"] * 10**4 // cnt).scaleb(-6)
            assert r["avg_price"] == decimal.Decimal(sums["sum_base_price"] * 10**4 // cnt).scaleb(-6)
            assert r["avg_disc"] == decimal.Decimal(int(disc[m].sum()) * 10**4 // cnt).scaleb(-6)
    assert seen_rows == int(keep.sum()) and len(got) == out.num_rows


def test_high_cardinality_30m_rows_against_torch(bb, ctx):
    """GROUP BY an Int64 key with 7.5 M distinct values over 30 M rows (the two-pass path: sketch, warp-synchronous
    intern, global atomics): group set, counts and Int64/Decimal128 sums equal torch's sort-based group-by."""
    import torch

    n, card = 30_000_000, 7_500_000
    g = torch.Generator(device="cuda").manual_seed(5)
    key = torch.randint(0, card, (n,), device="cuda", generator=g) * 32 + 1
    v = torch.randint(-1000, 1000, (n,), device="cuda", generator=g)
    dec = torch.zeros((n, 2), dtype=torch.int64, device="cuda")
    dec[:, 0] = torch.randint(-10**12, 10**12, (n,), device="cuda", generator=g)
    dec[:, 1] = dec[:, 0] >> 63  # sign extension: a valid little-endian i128
    schema = pa.schema([("k", pa.int64()), ("v", pa.int64()), ("d", pa.decimal128(15, 2))])
    batch = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=key.data_ptr()), dict(values=v.data_ptr()),
                                                     dict(values=dec.data_ptr())], n, keepalive=(key, v, dec))
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], [("sum", 1, "sv"), ("sum", 2, "sd"), ("count", None, "c")],
                              schema)
    agg.update(batch)
    assert agg.last_path == "global"
    out = agg.emit().download()
    uk, inv = torch.unique(key, return_inverse=True)
    assert out.num_rows == uk.numel()
    zero = torch.zeros(uk.numel(), dtype=torch.int64, device="cuda")
    want_v = zero.clone().index_add_(0, inv, v).cpu().numpy()
    want_d = zero.clone().index_add_(0, inv, dec[:, 0]).cpu().numpy()
    want_c = torch.bincount(inv, minlength=uk.numel()).cpu().numpy()
    gk = out.column(0).to_numpy()
    order = np.argsort(gk)
    assert np.array_equal(gk[order], uk.cpu().numpy())
    assert np.array_equal(out.column(1).to_numpy()[order], want_v)
    # Decimal128 sums: compare the low 64 bits and the sign-extended high word of the raw buffers
    dcol = out.column(2)
    raw = np.frombuffer(dcol.buffers()[1], dtype=np.int64, count=2 * len(dcol)).reshape(-1, 2)[order]
    assert np.array_equal(raw[:, 0], want_d) and np.array_equal(raw[:, 1], want_d >> 63)
    assert np.array_equal(out.column(out.num_columns - 1).to_numpy()[order], want_c)

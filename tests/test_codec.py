"""GpuPhysicalCodec (SURVEY.md §8b "Serialization"): the plan nodes of the two GPU operators survive
try_encode -> try_decode, the bytes are valid protobuf for proto/bgpu.proto (checked with google.protobuf's own runtime
on a descriptor built from the same field numbers), and a wrapping codec defers to its inner codec first
(/root/reference/ballista/tests/src/test_physical_codec.rs:29-64). No GPU needed."""
import datetime
import decimal
import os
import re
import sys

import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def proto_classes():
    """Message classes for proto/bgpu.proto built from the .proto text itself (a tiny parser for its flat syntax)."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    text = open(os.path.join(ROOT, "arrow-ballista_b200", "proto", "bgpu.proto")).read()
    text = re.sub(r"//[^\n]*", "", text)
    fd = descriptor_pb2.FileDescriptorProto(name="bgpu.proto", package="bgpu.protobuf", syntax="proto3")
    scalar = {"uint32": 13, "int32": 5, "uint64": 4, "fixed64": 6, "string": 9, "bytes": 12}
    for name, body in re.findall(r"message\s+(\w+)\s*\{((?:[^{}]|\{[^{}]*\})*)\}", text):
        m = fd.message_type.add(name=name)
        oneof = re.search(r"oneof\s+(\w+)\s*\{([^}]*)\}", body)
        if oneof:
            m.oneof_decl.add(name=oneof.group(1))
        for rep, typ, fname, num in re.findall(r"(repeated\s+)?(\w+)\s+(\w+)\s*=\s*(\d+)\s*;", body):
            f = m.field.add(name=fname, number=int(num), label=3 if rep else 1)
            if typ in scalar:
                f.type = scalar[typ]
            else:
                f.type, f.type_name = 11, f".bgpu.protobuf.{typ}"
            if oneof and re.search(rf"\b{fname}\b", oneof.group(2)):
                f.oneof_index = 0
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return {m.name: message_factory.GetMessageClass(pool.FindMessageTypeByName(f"bgpu.protobuf.{m.name}")) for m in fd.message_type}


def q1(bb):
    sys.path.insert(0, ROOT)
    import bench

    gi, aggs, filt = bench.q1_plan()
    return gi, aggs, filt, bench.lineitem_schema()


def test_shuffle_writer_node_round_trip(bb):
    schema = pa.schema([("a", pa.uint32()), ("b", pa.string())])

    class Child:
        def schema(self):
            return schema
    child = Child()
    node = bb.GpuShuffleWriterExec.try_new("jobOne", 3, [0, 5, 300], child, "/tmp/work", bb.Partitioning.Hash([1, 0], 16), None,
                                           input_schema=schema)
    codec = bb.GpuPhysicalCodec()
    buf = codec.try_encode(node)
    msg = proto_classes()["GpuPhysicalPlanNode"].FromString(buf)
    assert msg.WhichOneof("plan") == "shuffle_writer"
    sw = msg.shuffle_writer
    assert (sw.job_id, sw.stage_id, list(sw.partitions)) == ("jobOne", 3, [0, 5, 300])
    assert list(sw.output_partitioning.hash_expr) == [1, 0] and sw.output_partitioning.partition_count == 16
    back = codec.try_decode(buf, [child])
    assert isinstance(back, bb.GpuShuffleWriterExec)
    assert (back.job_id, back.stage_id, back.partitions, back.work_dir) == ("jobOne", 3, [0, 5, 300], "")  # serde/mod.rs:191
    assert back.shuffle_output_partitioning.exprs == [1, 0] and back.shuffle_output_partitioning.partition_count() == 16
    assert back.children() == [child] and back.schema().equals(schema)
    # un-partitioned stage output: no output_partitioning on the wire
    none = bb.GpuShuffleWriterExec.try_new("j", 1, [0], child, "", None, None, input_schema=schema)
    assert codec.try_decode(codec.try_encode(none), [child]).shuffle_output_partitioning is None
    with pytest.raises(bb.BgpuError):
        codec.try_decode(buf, [])  # the child rides in PhysicalExtensionNode.inputs and must be there


def test_aggregate_node_round_trip_q1(bb):
    gi, aggs, filt, schema = q1(bb)
    node = bb.GpuAggregateExecNode(bb.AggregateMode.Partial, gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs],
                                   schema, filt)
    codec = bb.GpuPhysicalCodec()
    buf = codec.try_encode(node)
    P = proto_classes()
    msg = P["GpuPhysicalPlanNode"].FromString(buf)
    assert msg.WhichOneof("plan") == "aggregate"
    a = msg.aggregate
    assert list(a.group_expr) == gi and a.mode == 0
    assert list(a.aggr_expr_name) == [n for _, _, n in aggs]
    assert list(a.group_expr_name) == ["l_returnflag", "l_linestatus"]
    assert list(a.aggr_fn) == [2, 2, 2, 2, 3, 3, 3, 4]
    assert len(a.aggr_expr[-1].nodes) == 0  # COUNT(*)
    assert [n.op for n in a.aggr_expr[2].nodes] == [0, 1, 0, 3, 4]  # price * (1 - disc), postfix
    assert len(a.filter.nodes) == 3
    assert pa.ipc.open_stream(pa.BufferReader(a.input_schema + b"\xff\xff\xff\xff\0\0\0\0")).schema.equals(schema)
    back = codec.try_decode(buf)
    assert isinstance(back, bb.GpuAggregateExecNode)
    assert back.mode == 0 and back.group_by == gi and back.input_schema.equals(schema)
    assert repr(back.filter) == repr(filt)
    assert [(f, repr(e), n) for f, e, n in back.aggr_expr] == [(bb.AggregateFunction.parse(f), repr(e), n) for f, e, n in aggs]
    assert codec.try_encode(back) == buf  # canonical: re-encoding the decoded node gives the same bytes
    # bytes written by google.protobuf for the same message decode to the same node (wire compatibility both ways)
    assert repr(codec.try_decode(msg.SerializeToString()).aggr_expr) == repr(back.aggr_expr)


def test_literal_kinds_survive(bb):
    from ballista_b200.expr import col, lit

    e = ((col(0) + lit(-7)) * lit(2.5)).and_(col(1) <= lit(datetime.date(1998, 9, 2))).and_(col(2) > lit(decimal.Decimal("-12.345")))
    node = bb.GpuAggregateExecNode(2, [0], [(2, e, "x")], pa.schema([("k", pa.int64()), ("d", pa.date32()), ("m", pa.decimal128(15, 3))]))
    back = bb.GpuPhysicalCodec().try_decode(bb.GpuPhysicalCodec().try_encode(node))
    got = back.aggr_expr[0][1]
    assert repr(got) == repr(e)
    assert [(n.op, n.a, n.precision, n.scale, n.lo, n.hi) for n in got.to_nodes()] == \
           [(n.op, n.a, n.precision, n.scale, n.lo, n.hi) for n in e.to_nodes()]


def test_wrapping_codec_tries_the_inner_codec_first(bb):
    class Inner:
        def try_encode(self, node):
            if node == "theirs":
                return b"INNER"
            raise ValueError("not mine")

        def try_decode(self, buf, inputs):
            if buf == b"INNER":
                return "theirs"
            raise ValueError("not mine")
    codec = bb.GpuPhysicalCodec(Inner())
    assert codec.try_encode("theirs") == b"INNER" and codec.try_decode(b"INNER") == "theirs"
    node = bb.GpuAggregateExecNode(0, [0], [(4, None, "c")], pa.schema([("k", pa.int64())]))
    assert codec.try_decode(codec.try_encode(node)).aggr_expr[0][2] == "c"
    with pytest.raises(bb.BgpuError):
        codec.try_encode(object())
    with pytest.raises(bb.BgpuError):
        codec.try_decode(b"\x1a\x00")  # field 3: not a GpuPhysicalPlanNode


@pytest.mark.gpu
def test_decoded_aggregate_runs_like_the_original(bb, ctx):
    """A stage plan that went through the codec computes the same Q1 Partial state as the operator built directly."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import lineitem_batch

    gi, aggs, filt, schema = q1(bb)
    direct = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, schema, filter=filt)
    codec = bb.GpuPhysicalCodec()
    decoded = codec.try_decode(codec.try_encode(direct), ctx=ctx)
    assert isinstance(decoded, bb.GpuAggregateExec)
    li = lineitem_batch(np.random.default_rng(4), 30_000)
    keys = [(n, "ascending") for n in ("l_returnflag", "l_linestatus")]
    a = pa.Table.from_batches([direct.execute([li])]).sort_by(keys)
    b = pa.Table.from_batches([decoded.execute([li])]).sort_by(keys)
    assert a.equals(b) and a.num_rows == 4


def test_mutated_plan_bytes_decode_or_fail_with_the_codecs_own_error(bb):
    """What arrives from the scheduler is bytes: flipped, truncated or spliced messages either decode or raise BgpuError
    (a task failure the executor reports), never another exception type."""
    import random

    gi, aggs, filt, schema = q1(bb)

    class Scan:
        def schema(self):
            return schema
    node = bb.GpuAggregateExecNode(0, gi, [(bb.AggregateFunction.parse(f), a, n) for f, a, n in aggs], schema, filt, inputs=[Scan()])
    writer = bb.GpuShuffleWriterExec.try_new("job", 1, [1, 2, 3], node, "", bb.Partitioning.Hash([0, 1], 16), None)
    codec = bb.GpuPhysicalCodec()
    goods = [codec.try_encode(node), codec.try_encode(writer)]
    R = random.Random(5)
    decoded = failed = 0
    for _ in range(1500):
        d = bytearray(R.choice(goods))
        k = R.randrange(3)
        if k == 0:
            for _ in range(R.randrange(1, 6)):
                d[R.randrange(len(d))] = R.randrange(256)
        elif k == 1:
            d = d[:R.randrange(len(d))]
        else:
            i = R.randrange(len(d))
            d[i:i + R.randrange(1, 5)] = bytes(R.randrange(256) for _ in range(R.randrange(0, 6)))
        try:
            codec.try_decode(bytes(d), [Scan()])
            decoded += 1
        except bb.BgpuError:
            failed += 1
    assert decoded + failed == 1500 and failed > 1000

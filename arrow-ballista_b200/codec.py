"""`GpuPhysicalCodec`: serialization of the two GPU plan nodes so that they ride in the scheduler's stage plans
(SURVEY.md §8b "Serialization").

Mirror of Ballista's `PhysicalExtensionCodec`
(/root/reference/ballista/core/src/serde/mod.rs:151-157 `try_decode(buf, inputs, registry)` / `try_encode(node, buf)`,
default implementation :274-398): the node's own fields are encoded as a protobuf message (schema in
`proto/bgpu.proto`, field numbers following `ShuffleWriterExecNode` /root/reference/ballista/core/proto/ballista.proto:47-55
and `AggregateExecNode` /root/reference/ballista/core/proto/datafusion.proto:1437-1450 where a counterpart exists), the
children ride next to it as `PhysicalExtensionNode.inputs` (/root/reference/ballista/core/proto/datafusion.proto:1136-1139).
A wrapping codec tries its inner codec first and falls back to its own messages
(/root/reference/ballista/tests/src/test_physical_codec.rs:29-64). As in the reference (serde/mod.rs:191) a decoded
shuffle writer has an empty work_dir: the executor re-creates it with its own (execution_engine.rs:74-93).

The protobuf wire format is written and read by hand (varints, length-delimited fields, fixed64): the messages are a
dozen fields and this keeps the module free of generated code. `tests/test_codec.py` parses the bytes with
google.protobuf's own runtime against the same schema.
"""
import struct
from dataclasses import dataclass, field

import pyarrow as pa

from . import ipc
from ._lib import BGPU_EINVAL, BgpuError
from .expr import BinaryExpr, Column, Literal, X_COLUMN, X_LITERAL, T_FLOAT64
from .partitioner import Partitioning

WT_VARINT, WT_FIXED64, WT_LEN = 0, 1, 2


# ---- protobuf wire format -----------------------------------------------------------------------------------------------
def _varint(v):
    v &= (1 << 64) - 1  # negative int32/int64 are sign-extended to 64 bits on the wire
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _tag(num, wt):
    return _varint((num << 3) | wt)


def _f_varint(num, v):
    return _tag(num, WT_VARINT) + _varint(int(v))


def _f_len(num, payload):
    return _tag(num, WT_LEN) + _varint(len(payload)) + bytes(payload)


def _f_fixed64(num, v):
    return _tag(num, WT_FIXED64) + struct.pack("<Q", v & ((1 << 64) - 1))


def _f_packed(num, values):
    return _f_len(num, b"".join(_varint(int(v)) for v in values)) if values else b""


def _fields(buf):
    """[(field number, wire type, value)] of one message; value = int (varint / fixed64) or bytes (length-delimited)."""
    out, i, n = [], 0, len(buf)

    def varint():
        nonlocal i
        v = shift = 0
        while True:
            if i >= n:
                raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: truncated varint")
            b = buf[i]
            i += 1
            v |= (b & 0x7F) << shift
            if not b & 0x80:
                return v
            shift += 7
            if shift > 63:
                raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: varint too long")
    while i < n:
        key = varint()
        num, wt = key >> 3, key & 7
        if wt == WT_VARINT:
            out.append((num, wt, varint()))
        elif wt == WT_FIXED64:
            if i + 8 > n:
                raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: truncated fixed64")
            out.append((num, wt, struct.unpack_from("<Q", buf, i)[0]))
            i += 8
        elif wt == WT_LEN:
            ln = varint()
            if i + ln > n:
                raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: truncated field")
            out.append((num, wt, bytes(buf[i:i + ln])))
            i += ln
        else:
            raise BgpuError(BGPU_EINVAL, f"GpuPhysicalCodec: unsupported wire type {wt}")
    return out


def _unpack_varints(payload):
    vals, j = [], 0
    while j < len(payload):
        x = shift = 0
        while True:
            if j >= len(payload):
                raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: truncated packed field")
            b = payload[j]
            j += 1
            x |= (b & 0x7F) << shift
            if not b & 0x80:
                break
            shift += 7
        vals.append(x)
    return vals


def _repeated_u32(fields, num):
    """A repeated scalar field, packed or not (a proto3 reader accepts both)."""
    out = []
    for f, wt, v in fields:
        if f == num:
            out.extend(_unpack_varints(v) if wt == WT_LEN else [v])
    return out


def _sint32(v):
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


# ---- messages ------------------------------------------------------------------------------------------------------------
def _encode_expr(e):
    """GpuExpr{repeated GpuExprNode nodes = 1}: the postfix bgpu_expr_node program (include/bgpu.h); None = COUNT(*)."""
    out = b""
    for n in ([] if e is None else e.to_nodes()):
        out += _f_len(1, _f_varint(1, n.op) + _f_varint(2, n.a) + _f_varint(3, n.precision) + _f_varint(4, n.scale)
                      + _f_fixed64(5, n.lo) + _f_fixed64(6, n.hi))
    return out


def _decode_expr(buf):
    stack = []
    for num, _, payload in _fields(buf):
        if num != 1:
            continue
        f = {k: v for k, _, v in _fields(payload)}
        op, a = f.get(1, 0), _sint32(f.get(2, 0))
        if op == X_COLUMN:
            stack.append(Column(a))
        elif op == X_LITERAL:
            lo, hi = f.get(5, 0), f.get(6, 0)
            if a == T_FLOAT64:
                value = struct.unpack("<d", struct.pack("<Q", lo))[0]
            else:
                value = (hi << 64) | lo
                if value >> 127:
                    value -= 1 << 128
            stack.append(Literal(value, a, _sint32(f.get(3, 0)), _sint32(f.get(4, 0))))
        else:
            if len(stack) < 2:
                raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: malformed expression program")
            r = stack.pop()
            l = stack.pop()
            stack.append(BinaryExpr(op, l, r))
    if len(stack) > 1:
        raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: malformed expression program")
    return stack[0] if stack else None


@dataclass
class GpuAggregateExecNode:
    """Decoded `GpuAggregateExecNode`; `instantiate(ctx)` builds the operator on a device context."""
    mode: int
    group_by: list
    aggr_expr: list  # [(fn, Expr | None, name)]
    input_schema: object
    filter: object = None
    group_expr_name: list = field(default_factory=list)
    inputs: list = field(default_factory=list)

    def children(self):
        return list(self.inputs)

    def with_new_children(self, children):
        import dataclasses

        return dataclasses.replace(self, inputs=list(children))

    def output_partitioning(self):
        # AggregateExec reports its input's partitioning (Partial) - hash-partitioned outputs are the exchange's business
        child = getattr(self.inputs[0], "output_partitioning", None) if self.inputs else None
        return child() if callable(child) else Partitioning.UnknownPartitioning(1)

    def execute(self, partition, ctx=None):
        """AggregateExec::execute(partition): the input's partition through one aggregate; yields its output batch."""
        from .runtime import Context

        agg = self.instantiate(ctx if ctx is not None else Context(0))
        try:
            for b in self.inputs[0].execute(partition):
                agg.update(b)
            yield agg.emit()
        finally:
            agg.close()

    def instantiate(self, ctx):
        from .aggregate import GpuAggregateExec

        return GpuAggregateExec(ctx, self.mode, self.group_by, self.aggr_expr, self.input_schema, filter=self.filter)


def _schema_bytes(schema):
    return ipc.schema_message(schema)


def _schema_from_bytes(b):
    r = ipc.StreamReader(ipc._HostContext(), b + ipc.EOS)
    try:
        return r.schema
    finally:
        r.close()


class GpuPhysicalCodec:
    """try_encode / try_decode for GpuShuffleWriterExec and GpuAggregateExec(-Node); anything else goes to `inner`."""

    def __init__(self, inner=None):
        self.inner = inner

    # -- encode ----------------------------------------------------------------------------------------------------------
    def try_encode(self, node):
        from .aggregate import GpuAggregateExec
        from .shuffle_writer import GpuShuffleWriterExec

        if self.inner is not None:
            try:
                return self.inner.try_encode(node)
            except Exception:
                pass  # not one of the inner codec's nodes: try ours (test_physical_codec.rs:49-63)
        if isinstance(node, GpuShuffleWriterExec):
            body = _f_len(1, node.job_id.encode()) + _f_varint(2, node.stage_id)
            part = node.shuffle_output_partitioning
            if part is not None:
                if part.kind != "hash":
                    raise BgpuError(BGPU_EINVAL, "Invalid shuffle partitioning scheme")
                body += _f_len(4, _f_packed(1, part.exprs) + _f_varint(2, part.partition_count()))
            body += _f_packed(5, node.partitions)
            return _f_len(1, body)
        if isinstance(node, (GpuAggregateExec, GpuAggregateExecNode)):
            body = _f_packed(1, node.group_by)
            for _, e, _ in node.aggr_expr:
                body += _f_len(2, _encode_expr(e))
            body += _f_varint(3, node.mode)
            for name in getattr(node, "group_expr_name", None) or [node.input_schema.field(i).name for i in node.group_by]:
                body += _f_len(5, name.encode())
            for _, _, name in node.aggr_expr:
                body += _f_len(6, name.encode())
            body += _f_len(7, _schema_bytes(node.input_schema))
            body += _f_packed(12, [f for f, _, _ in node.aggr_expr])
            flt = getattr(node, "filter", None)
            if flt is not None:
                body += _f_len(13, _encode_expr(flt))
            return _f_len(2, body)
        raise BgpuError(BGPU_EINVAL, f"GpuPhysicalCodec: unsupported plan node {type(node).__name__}")

    # -- decode ----------------------------------------------------------------------------------------------------------
    def try_decode(self, buf, inputs=(), ctx=None):
        """`inputs` = the already decoded children (PhysicalExtensionNode.inputs). A shuffle writer comes back as a
        GpuShuffleWriterExec with work_dir "" (serde/mod.rs:191); an aggregate as a GpuAggregateExecNode, or the
        operator itself when a device context is given."""
        from .shuffle_writer import GpuShuffleWriterExec

        if self.inner is not None:
            try:
                return self.inner.try_decode(buf, inputs)
            except Exception:
                pass
        try:
            return self._decode_own(buf, inputs, ctx)
        except BgpuError:
            raise
        except (AttributeError, TypeError, ValueError, LookupError, OverflowError, struct.error, pa.ArrowException) as e:
            # a field with the wrong wire type, a string that is not UTF-8, a schema Arrow refuses ...
            raise BgpuError(BGPU_EINVAL, f"GpuPhysicalCodec: malformed message ({type(e).__name__}: {e})") from e

    def _decode_own(self, buf, inputs, ctx):
        from .shuffle_writer import GpuShuffleWriterExec

        top = _fields(buf)
        if len(top) != 1 or top[0][1] != WT_LEN or top[0][0] not in (1, 2):
            raise BgpuError(BGPU_EINVAL, "GpuPhysicalCodec: not a GpuPhysicalPlanNode")
        which, _, body = top[0]
        f = _fields(body)
        if which == 1:
            if len(inputs) != 1:
                raise BgpuError(BGPU_EINVAL, "GpuShuffleWriterExecNode expects exactly one input")
            d = {k: v for k, _, v in f}
            part = None
            if 4 in d:
                pf = _fields(d[4])
                part = Partitioning.Hash(_repeated_u32(pf, 1), {k: v for k, _, v in pf}.get(2, 0))
            child = inputs[0]
            schema = child.schema() if callable(getattr(child, "schema", None)) else getattr(child, "schema", None)
            return GpuShuffleWriterExec.try_new(d.get(1, b"").decode(), d.get(2, 0), _repeated_u32(f, 5), child, "", part, ctx,
                                                input_schema=schema)
        d = {k: v for k, _, v in f}
        exprs = [_decode_expr(v) for k, _, v in f if k == 2]
        names = [v.decode() for k, _, v in f if k == 6]
        fns = _repeated_u32(f, 12)
        if not (len(exprs) == len(names) == len(fns)) or 7 not in d:
            raise BgpuError(BGPU_EINVAL, "GpuAggregateExecNode: aggr_expr / aggr_expr_name / aggr_fn / input_schema mismatch")
        node = GpuAggregateExecNode(d.get(3, 0), _repeated_u32(f, 1), list(zip(fns, exprs, names)), _schema_from_bytes(d[7]),
                                    _decode_expr(d[13]) if 13 in d else None, [v.decode() for k, _, v in f if k == 5], list(inputs))
        return node.instantiate(ctx) if ctx is not None else node

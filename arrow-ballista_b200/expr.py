"""Physical expressions accepted by GpuAggregateExec for aggregate arguments and the fused filter.

Mirrors the subset of DataFusion's `PhysicalExpr` tree that the hot path's plans carry (Column, Literal,
BinaryExpr with + - * and comparisons / AND): TPC-H Q1's `sum(l_extendedprice * (1 - l_discount))`
(/root/reference/benchmarks/queries/q1.sql) is `col(5) * (lit(1) - col(6))`. `to_nodes()` flattens a tree into the
postfix `bgpu_expr_node` program of include/bgpu.h.
"""
import datetime
import decimal
import struct

from ._lib import ExprNode

X_COLUMN, X_LITERAL, X_ADD, X_SUB, X_MUL, X_LT, X_LE, X_GT, X_GE, X_EQ, X_NE, X_AND = range(12)
T_INT64, T_FLOAT64, T_DECIMAL128, T_DATE32 = range(4)
_OP_NAMES = {X_ADD: "+", X_SUB: "-", X_MUL: "*", X_LT: "<", X_LE: "<=", X_GT: ">", X_GE: ">=", X_EQ: "=", X_NE: "!=",
             X_AND: "AND"}


class Expr:
    def _bin(self, op, other, swap=False):
        other = other if isinstance(other, Expr) else lit(other)
        return BinaryExpr(op, other, self) if swap else BinaryExpr(op, self, other)

    def __add__(self, o): return self._bin(X_ADD, o)
    def __radd__(self, o): return self._bin(X_ADD, o, True)
    def __sub__(self, o): return self._bin(X_SUB, o)
    def __rsub__(self, o): return self._bin(X_SUB, o, True)
    def __mul__(self, o): return self._bin(X_MUL, o)
    def __rmul__(self, o): return self._bin(X_MUL, o, True)
    def __lt__(self, o): return self._bin(X_LT, o)
    def __le__(self, o): return self._bin(X_LE, o)
    def __gt__(self, o): return self._bin(X_GT, o)
    def __ge__(self, o): return self._bin(X_GE, o)
    def eq(self, o): return self._bin(X_EQ, o)
    def ne(self, o): return self._bin(X_NE, o)
    def and_(self, o): return self._bin(X_AND, o)

    def to_nodes(self):
        out = []
        self._emit(out)
        return out


class Column(Expr):
    def __init__(self, index):
        self.index = int(index)

    def _emit(self, out):
        out.append(ExprNode(X_COLUMN, self.index, 0, 0, 0, 0))

    def __repr__(self):
        return f"#{self.index}"


class Literal(Expr):
    """ScalarValue: Int64, Float64, Decimal128(p, s) or Date32."""

    def __init__(self, value, type_=None, precision=0, scale=0):
        if type_ is None:
            if isinstance(value, bool):
                raise TypeError("boolean literals are not supported")
            if isinstance(value, int):
                type_ = T_INT64
            elif isinstance(value, float):
                type_ = T_FLOAT64
            elif isinstance(value, decimal.Decimal):
                type_ = T_DECIMAL128
                sign, digits, exp = value.as_tuple()
                scale = max(0, -exp)
                unscaled = int(value.scaleb(scale))
                precision = max(len(digits) + max(exp, 0), scale, 1)
                value = unscaled
            elif isinstance(value, datetime.date):
                type_ = T_DATE32
                value = (value - datetime.date(1970, 1, 1)).days
            else:
                raise TypeError(f"unsupported literal {value!r}")
        self.value, self.type, self.precision, self.scale = value, type_, int(precision), int(scale)

    def _emit(self, out):
        if self.type == T_FLOAT64:
            lo, hi = struct.unpack("<Q", struct.pack("<d", float(self.value)))[0], 0
        else:
            v = int(self.value) & ((1 << 128) - 1)
            lo, hi = v & ((1 << 64) - 1), v >> 64
        out.append(ExprNode(X_LITERAL, self.type, self.precision, self.scale, lo, hi))

    def __repr__(self):
        return f"lit({self.value})"


class BinaryExpr(Expr):
    def __init__(self, op, left, right):
        self.op, self.left, self.right = op, left, right

    def _emit(self, out):
        self.left._emit(out)
        self.right._emit(out)
        out.append(ExprNode(self.op, 0, 0, 0, 0, 0))

    def __repr__(self):
        return f"({self.left!r} {_OP_NAMES[self.op]} {self.right!r})"


def col(index):
    return Column(index)


def lit(value, type_=None, precision=0, scale=0):
    return value if isinstance(value, Expr) else Literal(value, type_, precision, scale)

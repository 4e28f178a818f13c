// Host-side compiler: postfix expression programs (include/bgpu.h) -> HValue / HFilterTerm (agg_plan.h).
// The accepted shapes are products of up to three `x`, `x + y`, `x - y` factors over columns and literals (what
// TPC-H Q1 and the db-benchmark group-by queries feed their aggregates) and conjunctions of `column <cmp> literal`.
// Anything else is BGPU_EUNSUPPORTED: there is no slower generic path hiding behind this one.
#include <string.h>

#include "agg_plan.h"

namespace bgpu {

ValKind vk_of(const Column& c) {
  if (c.kind != K_FIXED) return V_NONE;
  switch (c.format[0]) {
    case 'c': return V_I8;  case 'C': return V_U8;  case 's': return V_I16; case 'S': return V_U16;
    case 'i': return V_I32; case 'I': return V_U32; case 'l': return V_I64; case 'L': return V_U64;
    case 'f': return V_F32; case 'g': return V_F64; case 'd': return V_I128;
    case 't':
      if (c.format[1] == 'd') return c.format[2] == 'D' ? V_I32 : V_I64;  // date32 / date64
      return c.width == 4 ? V_I32 : V_I64;                                  // time / timestamp / duration
    default: return V_NONE;
  }
}

namespace {
struct Node {
  int32_t op = 0;
  HAtom atom;
  int l = -1, r = -1;
};

bool is_float_vk(int vk) { return vk == V_F32 || vk == V_F64; }

int int_precision(int vk) {  // DataFusion's coercion of integers next to decimals
  switch (vk) {
    case V_I8: case V_U8: return 3;
    case V_I16: case V_U16: return 5;
    case V_I32: case V_U32: return 10;
    default: return 20;
  }
}

int parse_tree(bgpu_ctx* ctx, const std::vector<Column>& fields, const bgpu_expr_node* nodes, int32_t n,
               std::vector<Node>* out, int* root) {
  std::vector<int> stack;
  for (int32_t i = 0; i < n; i++) {
    const bgpu_expr_node& e = nodes[i];
    Node nd;
    nd.op = e.op;
    if (e.op == BGPU_X_COLUMN) {
      if (e.a < 0 || e.a >= (int32_t)fields.size()) return ctx->fail(BGPU_EINVAL, "expression column out of range");
      const Column& f = fields[(size_t)e.a];
      const ValKind vk = vk_of(f);
      if (vk == V_NONE) return ctx->fail(BGPU_EUNSUPPORTED, "expression over non-numeric column " + f.name);
      nd.atom.is_col = true;
      nd.atom.col = e.a;
      nd.atom.vk = vk;
      if (vk == V_I128) {
        nd.atom.cls = CLS_I128;
        sscanf(f.format.c_str(), "d:%d,%d", &nd.atom.precision, &nd.atom.scale);
      } else if (is_float_vk(vk)) {
        nd.atom.cls = CLS_F64;
      } else {
        nd.atom.cls = CLS_I64;
        nd.atom.precision = int_precision(vk);
      }
    } else if (e.op == BGPU_X_LITERAL) {
      nd.atom.is_col = false;
      const __int128 v = (__int128)(((unsigned __int128)e.hi << 64) | e.lo);
      switch (e.a) {
        case BGPU_T_INT64: nd.atom.cls = CLS_I64; nd.atom.lit = (int64_t)e.lo; nd.atom.precision = 20; break;
        case BGPU_T_DATE32: nd.atom.cls = CLS_I64; nd.atom.lit = (int32_t)e.lo; nd.atom.precision = 10; break;
        case BGPU_T_DECIMAL128:
          nd.atom.cls = CLS_I128; nd.atom.lit = v; nd.atom.precision = e.precision; nd.atom.scale = e.scale;
          break;
        case BGPU_T_FLOAT64: {
          nd.atom.cls = CLS_F64;
          double d;
          memcpy(&d, &e.lo, 8);
          nd.atom.flit = d;
        } break;
        default: return ctx->fail(BGPU_EINVAL, "bad literal type");
      }
    } else if (e.op >= BGPU_X_ADD && e.op <= BGPU_X_AND) {
      if (stack.size() < 2) return ctx->fail(BGPU_EINVAL, "malformed postfix expression");
      nd.r = stack.back(); stack.pop_back();
      nd.l = stack.back(); stack.pop_back();
    } else {
      return ctx->fail(BGPU_EINVAL, "unknown expression op");
    }
    out->push_back(nd);
    stack.push_back((int)out->size() - 1);
  }
  if (stack.size() != 1) return ctx->fail(BGPU_EINVAL, "malformed postfix expression");
  *root = stack[0];
  return BGPU_OK;
}

bool is_atom(const Node& n) { return n.op == BGPU_X_COLUMN || n.op == BGPU_X_LITERAL; }

bool pow10_i128(int k, __int128* out) {
  if (k < 0 || k > 38) return false;
  __int128 p = 1;
  for (int i = 0; i < k; i++) p *= 10;
  *out = p;
  return true;
}
}  // namespace

int compile_value(bgpu_ctx* ctx, const std::vector<Column>& fields, const bgpu_expr_node* nodes, int32_t n,
                  HValue* out) {
  *out = HValue();
  if (n == 0) return BGPU_OK;  // COUNT(*)
  std::vector<Node> t;
  int root = -1;
  BGPU_TRY(parse_tree(ctx, fields, nodes, n, &t, &root));
  // left-deep product of factors
  std::vector<int> factors;
  int cur = root;
  while (t[(size_t)cur].op == BGPU_X_MUL) {
    if (t[(size_t)t[(size_t)cur].r].op == BGPU_X_MUL)
      return ctx->fail(BGPU_EUNSUPPORTED, "aggregate argument: only left-deep products are supported");
    factors.insert(factors.begin(), t[(size_t)cur].r);
    cur = t[(size_t)cur].l;
  }
  factors.insert(factors.begin(), cur);
  if ((int)factors.size() > kMaxFactors)
    return ctx->fail(BGPU_EUNSUPPORTED, "aggregate argument: more than 3 factors");
  int32_t cls = CLS_I64;
  bool any_op = factors.size() > 1;
  for (int fi : factors) {
    const Node& fn = t[(size_t)fi];
    HFactor f;
    if (is_atom(fn)) {
      f.x = fn.atom;
    } else if (fn.op == BGPU_X_ADD || fn.op == BGPU_X_SUB) {
      if (!is_atom(t[(size_t)fn.l]) || !is_atom(t[(size_t)fn.r]))
        return ctx->fail(BGPU_EUNSUPPORTED, "aggregate argument: nested sums are not supported");
      f.x = t[(size_t)fn.l].atom;
      f.y = t[(size_t)fn.r].atom;
      f.op = fn.op == BGPU_X_ADD ? 1 : 2;
      any_op = true;
    } else {
      return ctx->fail(BGPU_EUNSUPPORTED, "aggregate argument: unsupported operator");
    }
    for (const HAtom* a : {&f.x, &f.y}) {
      if (a == &f.y && f.op == 0) continue;
      if (a->cls == CLS_I128) { if (cls == CLS_F64) cls = -1; else cls = CLS_I128; }
      if (a->cls == CLS_F64) { if (cls == CLS_I128) cls = -1; else cls = CLS_F64; }
    }
    out->f.push_back(f);
  }
  if (cls < 0) return ctx->fail(BGPU_EUNSUPPORTED, "aggregate argument mixes Decimal128 and Float64");
  out->cls = cls;
  // type rules
  for (HFactor& f : out->f) {
    for (HAtom* a : {&f.x, &f.y}) {
      if (a == &f.y && f.op == 0) continue;
      const int32_t vk = a->vk;
      if (cls == CLS_F64) {
        if (a->is_col ? (any_op && vk != V_F64) : (a->cls != CLS_F64 && a->cls != CLS_I64))
          return ctx->fail(BGPU_EUNSUPPORTED, "Float64 arithmetic needs Float64 columns");
        if (!a->is_col && a->cls == CLS_I64) a->flit = (double)(int64_t)a->lit;  // literal coerced to Float64
      } else if (cls == CLS_I64) {
        if (a->is_col && any_op && vk != V_I64 && vk != V_U64)
          return ctx->fail(BGPU_EUNSUPPORTED, "integer arithmetic needs Int64 columns");
      }  // CLS_I128: integers are coerced to Decimal128(int_precision, 0): precision/scale already set
    }
    if (cls == CLS_I128) {
      const HAtom& x = f.x;
      if (f.op == 0) {
        f.precision = x.precision; f.scale = x.scale;
      } else {
        const HAtom& y = f.y;
        f.scale = std::max(x.scale, y.scale);
        f.precision = std::min(38, f.scale + std::max(x.precision - x.scale, y.precision - y.scale) + 1);
        if (!pow10_i128(f.scale - x.scale, &f.x_mul) || !pow10_i128(f.scale - y.scale, &f.y_mul))
          return ctx->fail(BGPU_EUNSUPPORTED, "decimal rescale out of range");
      }
    }
  }
  if (cls == CLS_I128) {
    int p = out->f[0].precision, s = out->f[0].scale;
    for (size_t i = 1; i < out->f.size(); i++) {
      p = std::min(38, p + out->f[i].precision + 1);
      s += out->f[i].scale;
    }
    if (s > 38) return ctx->fail(BGPU_EUNSUPPORTED, "Output scale of decimal multiply would exceed 38");
    out->precision = p;
    out->scale = s;
  }
  if (out->plain_column()) out->vk = out->f[0].x.vk;
  else out->vk = cls == CLS_I128 ? V_I128 : cls == CLS_F64 ? V_F64 : V_I64;
  return BGPU_OK;
}

static bool atoms_equal(const HAtom& a, const HAtom& b) {
  if (a.is_col != b.is_col) return false;
  if (a.is_col) return a.col == b.col;
  return a.cls == b.cls && a.lit == b.lit && a.flit == b.flit && a.precision == b.precision && a.scale == b.scale;
}
static bool factors_equal(const HFactor& a, const HFactor& b) {
  return a.op == b.op && atoms_equal(a.x, b.x) && (a.op == 0 || atoms_equal(a.y, b.y));
}
bool values_equal(const HValue& a, const HValue& b) {
  if (a.cls != b.cls || a.f.size() != b.f.size()) return false;
  for (size_t i = 0; i < a.f.size(); i++)
    if (!factors_equal(a.f[i], b.f[i])) return false;
  return true;
}
bool value_extends(const HValue& a, const HValue& b) {
  if (a.cls != b.cls || a.f.empty() || b.f.size() != a.f.size() + 1) return false;
  for (size_t i = 0; i < a.f.size(); i++)
    if (!factors_equal(a.f[i], b.f[i])) return false;
  return true;
}

int compile_filter(bgpu_ctx* ctx, const std::vector<Column>& fields, const bgpu_expr_node* nodes, int32_t n,
                   std::vector<HFilterTerm>* out) {
  out->clear();
  if (n == 0) return BGPU_OK;
  std::vector<Node> t;
  int root = -1;
  BGPU_TRY(parse_tree(ctx, fields, nodes, n, &t, &root));
  std::vector<int> todo{root};
  while (!todo.empty()) {
    const int i = todo.back();
    todo.pop_back();
    const Node& nd = t[(size_t)i];
    if (nd.op == BGPU_X_AND) {
      todo.push_back(nd.r);
      todo.push_back(nd.l);
      continue;
    }
    if (nd.op < BGPU_X_LT || nd.op > BGPU_X_NE) return ctx->fail(BGPU_EUNSUPPORTED, "filter: expected a comparison");
    const Node *l = &t[(size_t)nd.l], *r = &t[(size_t)nd.r];
    int cmp = nd.op;
    if (!is_atom(*l) || !is_atom(*r)) return ctx->fail(BGPU_EUNSUPPORTED, "filter: only column <cmp> literal");
    if (!l->atom.is_col) {  // literal <cmp> column  ->  column <flipped cmp> literal
      std::swap(l, r);
      cmp = cmp == BGPU_X_LT ? BGPU_X_GT : cmp == BGPU_X_LE ? BGPU_X_GE : cmp == BGPU_X_GT ? BGPU_X_LT
            : cmp == BGPU_X_GE ? BGPU_X_LE : cmp;
    }
    if (!l->atom.is_col || r->atom.is_col) return ctx->fail(BGPU_EUNSUPPORTED, "filter: only column <cmp> literal");
    HFilterTerm ft;
    ft.col = l->atom.col;
    ft.cmp = cmp;
    ft.vk = vk_of(fields[(size_t)ft.col]);
    const HAtom& lit = r->atom;
    if (l->atom.cls == CLS_F64) {
      if (lit.cls == CLS_I128) return ctx->fail(BGPU_EUNSUPPORTED, "filter: Float64 column against a decimal literal");
      ft.flit = lit.cls == CLS_F64 ? lit.flit : (double)(int64_t)lit.lit;
    } else if (l->atom.cls == CLS_I128) {
      if (lit.cls == CLS_F64 || lit.scale > l->atom.scale)
        return ctx->fail(BGPU_EUNSUPPORTED, "filter: literal cannot be rescaled to the column's decimal type");
      __int128 m;
      if (!pow10_i128(l->atom.scale - lit.scale, &m) || __builtin_mul_overflow(lit.lit, m, &ft.lit))
        return ctx->fail(BGPU_EUNSUPPORTED, "filter: literal out of range");
    } else {
      if (lit.cls != CLS_I64) return ctx->fail(BGPU_EUNSUPPORTED, "filter: integer column needs an integer literal");
      ft.lit = lit.lit;
    }
    out->push_back(ft);
  }
  if ((int)out->size() > kMaxFilterTerms) return ctx->fail(BGPU_EUNSUPPORTED, "filter: more than 4 conjuncts");
  return BGPU_OK;
}

}  // namespace bgpu

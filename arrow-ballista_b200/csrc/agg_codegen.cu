// Generates the plan-specific `struct Plan` that agg_jit_prelude.cuh's kernel templates are instantiated over:
// straight-line CUDA for the row filter, the packed group key and every aggregate argument of ONE aggregate bound to
// ONE batch shape (which columns carry validity, which dictionaries need a remap). No interpretation is left for the
// device: types, literals, decimal rescale factors, key byte offsets and accumulator slots are compile-time constants.
//
// Reference semantics restated (none of it is in /root/reference): see agg_plan.h.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "agg_host.h"

using namespace bgpu;

static const char kPrelude[] =
#include "agg_jit_prelude.inc"
    ;

const char* agg_jit_prelude() { return kPrelude; }

int32_t jit_slot(JitShape* sh, int32_t batch_col) {
  for (size_t i = 0; i < sh->slots.size(); i++)
    if (sh->slots[i].batch_col == batch_col) return (int32_t)i;
  if ((int)sh->slots.size() >= kJitSlots) return -1;
  JitSlot s;
  s.batch_col = batch_col;
  sh->slots.push_back(s);
  return (int32_t)sh->slots.size() - 1;
}

namespace {

struct Out {
  std::string s;
  void line(const std::string& l) { s += l; s += '\n'; }
  void fmt(const char* f, ...) __attribute__((format(printf, 2, 3))) {
    char buf[1024];
    va_list ap;
    va_start(ap, f);
    vsnprintf(buf, sizeof buf, f, ap);
    va_end(ap);
    s += buf;
    s += '\n';
  }
};

std::string hex64(uint64_t v) {
  char b[32];
  snprintf(b, sizeof b, "0x%llxULL", (unsigned long long)v);
  return b;
}
std::string val_lit(__int128 v) {
  return "Val{" + hex64((uint64_t)v) + ", " + hex64((uint64_t)((unsigned __int128)v >> 64)) + "}";
}
std::string f64_lit(double d) {
  uint64_t b;
  memcpy(&b, &d, 8);
  return "as_f64(" + hex64(b) + ")";
}

const char* vk_ctype(int vk) {
  switch (vk) {
    case V_I8: return "int8_t";   case V_U8: return "uint8_t";
    case V_I16: return "int16_t"; case V_U16: return "uint16_t";
    case V_I32: return "int32_t"; case V_U32: return "uint32_t";
    case V_I64: return "int64_t"; case V_U64: return "uint64_t";
    case V_F32: return "float";   case V_F64: return "double";
    default: return "Val";
  }
}
bool vk_signed(int vk) { return vk == V_I8 || vk == V_I16 || vk == V_I32 || vk == V_I64; }
bool vk_float(int vk) { return vk == V_F32 || vk == V_F64; }
const char* width_utype(int w) {
  switch (w) {
    case 1: return "uint8_t";
    case 2: return "uint16_t";
    case 4: return "uint32_t";
    case 8: return "uint64_t";
    default: return "Val";
  }
}

// `x.c<s>` as an i128 Val
std::string as_i128(int s, int vk) {
  const std::string c = "x.c" + std::to_string(s);
  if (vk == V_I128) return c;
  return vk_signed(vk) ? "sext((int64_t)" + c + ")" : "zext((uint64_t)" + c + ")";
}
std::string as_double(int s, int vk) {
  const std::string c = "x.c" + std::to_string(s);
  if (vk == V_F64) return c;
  if (vk == V_F32) return "(double)" + c;
  return vk_signed(vk) ? "(double)(int64_t)" + c : "(double)(uint64_t)" + c;
}
std::string as_u64(int s, int vk) {  // Int64 class: sign-/zero-extended
  const std::string c = "x.c" + std::to_string(s);
  return vk_signed(vk) ? "(uint64_t)(int64_t)" + c : "(uint64_t)" + c;
}

struct Gen {
  const bgpu_agg* g;
  const JitShape& sh;
  Out o;
  int slot_of(int32_t batch_col) const {
    for (size_t i = 0; i < sh.slots.size(); i++)
      if (sh.slots[i].batch_col == batch_col) return (int)i;
    return -1;
  }

  // "early" slots feed the addresses of other loads (utf8 offsets; dictionary keys that go through a remap): they are
  // loaded one tile ahead (load0 / Raw0) so that the dependent loads can be issued together with everything else
  static bool early(const JitSlot& sl) {
    return sl.key_use && (sl.kind == K_UTF8 || sl.kind == K_BINARY || (sl.kind == K_DICT32 && sl.remap));
  }

  void raw_struct() {
    o.line("  struct Raw0 {");
    for (size_t s = 0; s < sh.slots.size(); s++) {
      const JitSlot& sl = sh.slots[s];
      if (!early(sl)) continue;
      if (sl.kind == K_DICT32) {
        if (sl.nullable) o.fmt("    bool v%zu;", s);
        o.fmt("    int32_t k%zu;", s);
      } else {
        o.fmt("    int32_t b%zu, l%zu;", s, s);
      }
    }
    o.line("  };");
    o.line("  struct Raw : Raw0 {");
    for (size_t s = 0; s < sh.slots.size(); s++) {
      const JitSlot& sl = sh.slots[s];
      if (sl.nullable && !(early(sl) && sl.kind == K_DICT32)) o.fmt("    bool v%zu;", s);
      if (sl.value_use || sl.filter_use) o.fmt("    %s c%zu;", vk_ctype(sl.vk), s);
      if (sl.key_use) {
        switch (sl.kind) {
          case K_FIXED: o.fmt("    %s k%zu;", width_utype(sl.width), s); break;
          case K_BOOL: o.fmt("    bool k%zu;", s); break;
          case K_DICT32: if (!early(sl)) o.fmt("    int32_t k%zu;", s); break;
          default: o.fmt("    StrWords w%zu;", s); break;
        }
      }
    }
    o.line("  };");
  }

  static std::string need_of(const JitSlot& sl, bool value_side) {
    // value_side: the typed c<s> field (filter | values); otherwise the key field
    std::string n;
    if (value_side) {
      if (sl.filter_use) n += "NEED_FILTER";
      if (sl.value_use) n += std::string(n.empty() ? "" : " | ") + "NEED_VALUES";
    } else {
      n = "NEED_KEY";
    }
    return n;
  }

  void load0() {
    o.line("  template <int NEED> static __device__ __forceinline__ void load0(const JitArgs& a, Row r, Raw0& x) {");
    for (size_t s = 0; s < sh.slots.size(); s++) {
      const JitSlot& sl = sh.slots[s];
      if (!early(sl)) continue;
      if (sl.kind == K_DICT32) {
        if (sl.nullable) o.fmt("    if (NEED & NEED_KEY) x.v%zu = bit_set(a.valid[%zu], r);", s, s);
        o.fmt("    if (NEED & NEED_KEY) x.k%zu = __ldg((const int32_t*)a.col[%zu] + r);", s, s);
      } else {
        o.fmt("    if (NEED & NEED_KEY) { x.b%zu = __ldg(a.aux[%zu] + r); x.l%zu = __ldg(a.aux[%zu] + r + 1) - x.b%zu; }", s, s, s, s, s);
      }
    }
    o.line("  }");
  }

  // bulk L2 prefetch of rows [row0, row0 + n) of every fixed-width buffer the kernels stream (one thread per buffer)
  void prefetch() {
    o.line("  template <int NEED> static __device__ __forceinline__ void prefetch_l2(const JitArgs& a, int64_t row0, int64_t n) {");
    int lane = 0;
    for (size_t s = 0; s < sh.slots.size(); s++) {
      const JitSlot& sl = sh.slots[s];
      std::string any;
      if (sl.filter_use) any += "NEED_FILTER";
      if (sl.value_use) any += std::string(any.empty() ? "" : " | ") + "NEED_VALUES";
      if (sl.key_use) any += std::string(any.empty() ? "" : " | ") + "NEED_KEY";
      if (sl.kind == K_FIXED || sl.kind == K_DICT32)
        o.fmt("    if ((NEED & (%s)) && threadIdx.x == %d) bulk_prefetch_l2((const char*)a.col[%zu] + row0 * %d, n * %d);", any.c_str(),
              lane++, s, sl.kind == K_DICT32 ? 4 : sl.width, sl.kind == K_DICT32 ? 4 : sl.width);
      else if (sl.kind == K_UTF8 || sl.kind == K_BINARY)
        o.fmt("    if ((NEED & NEED_KEY) && threadIdx.x == %d) bulk_prefetch_l2((const char*)(a.aux[%zu] + row0), (n + 1) * 4);", lane++, s);
    }
    o.line("  }");
  }

  void load1() {
    o.line("  template <int NEED> static __device__ __forceinline__ void load1(const JitArgs& a, Row r, Raw& x) {");
    for (size_t s = 0; s < sh.slots.size(); s++) {
      const JitSlot& sl = sh.slots[s];
      std::string any;
      if (sl.filter_use) any += "NEED_FILTER";
      if (sl.value_use) any += std::string(any.empty() ? "" : " | ") + "NEED_VALUES";
      if (sl.key_use) any += std::string(any.empty() ? "" : " | ") + "NEED_KEY";
      const bool dict_early = early(sl) && sl.kind == K_DICT32;
      if (sl.nullable && !dict_early) o.fmt("    if (NEED & (%s)) x.v%zu = bit_set(a.valid[%zu], r);", any.c_str(), s, s);
      if (sl.value_use || sl.filter_use) {
        const std::string need = need_of(sl, true);
        if (sl.vk == V_I128) o.fmt("    if (NEED & (%s)) x.c%zu = ld_i128(a.col[%zu], r);", need.c_str(), s, s);
        else o.fmt("    if (NEED & (%s)) x.c%zu = __ldg((const %s*)a.col[%zu] + r);", need.c_str(), s, vk_ctype(sl.vk), s);
      }
      if (sl.key_use && !early(sl)) {
        switch (sl.kind) {
          case K_FIXED:
            if (sl.width == 16) o.fmt("    if (NEED & NEED_KEY) x.k%zu = ld_i128(a.col[%zu], r);", s, s);
            else o.fmt("    if (NEED & NEED_KEY) x.k%zu = __ldg((const %s*)a.col[%zu] + r);", s, width_utype(sl.width), s);
            break;
          case K_BOOL: o.fmt("    if (NEED & NEED_KEY) x.k%zu = bit_set((const uint8_t*)a.col[%zu], r);", s, s); break;
          default: o.fmt("    if (NEED & NEED_KEY) x.k%zu = __ldg((const int32_t*)a.col[%zu] + r);", s, s); break;
        }
      }
    }
    o.line("  }");
  }

  void load2() {
    o.line("  template <int NEED> static __device__ __forceinline__ void load2(const JitArgs& a, Row r, Raw& x) {");
    for (size_t ki = 0; ki < sh.key_slot.size(); ki++) {
      const size_t s = (size_t)sh.key_slot[ki];
      const JitSlot& sl = sh.slots[s];
      const int width = g->key_width[ki];
      const std::string guard = sl.nullable ? " && x.v" + std::to_string(s) : "";
      if (sl.kind == K_DICT32 && sl.remap) {
        o.fmt("    if ((NEED & NEED_KEY)%s) x.k%zu = __ldg(a.aux[%zu] + x.k%zu);", guard.c_str(), s, s, s);
      } else if (sl.kind == K_UTF8 || sl.kind == K_BINARY) {
        // the words holding the string's bytes (assembled into the key field by key())
        // (offsets are valid for null rows too: no need to wait for the validity bit)
        o.fmt("    if (NEED & NEED_KEY) x.w%zu = str_load<%d>(a.col[%zu], x.b%zu, x.l%zu);", s, width, s, s, s);
      }
    }
    o.line("  }");
  }

  void filter() {
    o.line("  static __device__ __forceinline__ bool filter(const JitArgs& a, const Raw& x) {");
    for (size_t i = 0; i < g->filter.size(); i++) {
      const HFilterTerm& t = g->filter[i];
      const size_t s = (size_t)sh.filter_slot[i];
      const JitSlot& sl = sh.slots[s];
      if (sl.nullable) o.fmt("    if (!x.v%zu) return false;", s);
      o.line("    {");
      if (vk_float(t.vk)) {
        o.fmt("      const double v = %s;", as_double((int)s, t.vk).c_str());
        o.line("      if (v != v) return false;  // NaN compares false (NaN filters are off the path)");
        o.fmt("      const int c = f64_cmp(v, %s);", f64_lit(t.flit).c_str());
      } else {
        o.fmt("      const int c = i128_cmp(%s, %s);", as_i128((int)s, t.vk).c_str(), val_lit(t.lit).c_str());
      }
      const char* rel = t.cmp == BGPU_X_LT ? "c < 0" : t.cmp == BGPU_X_LE ? "c <= 0" : t.cmp == BGPU_X_GT ? "c > 0"
                        : t.cmp == BGPU_X_GE ? "c >= 0" : t.cmp == BGPU_X_EQ ? "c == 0" : "c != 0";
      o.fmt("      if (!(%s)) return false;", rel);
      o.line("    }");
    }
    o.line("    return true;");
    o.line("  }");
  }

  void key() {
    o.line("  static __device__ __forceinline__ PackedKey key(const JitArgs& a, const Raw& x) {");
    o.line("    PackedKey k{0, 0, 0};");
    o.line("    bool too_long = false;  // a string key beyond its packed width: the pass is void (flags[1])");
    for (size_t ki = 0; ki < sh.key_slot.size(); ki++) {
      const size_t s = (size_t)sh.key_slot[ki];
      const JitSlot& sl = sh.slots[s];
      const int off = g->key_off[ki], width = g->key_width[ki];
      const std::string word = off < 8 ? "k.k0" : "k.k1";
      const int shift = 8 * (off < 8 ? off : off - 8);
      std::string body;
      char buf[512];
      switch (sl.kind) {
        case K_FIXED:
          if (sl.width == 16) snprintf(buf, sizeof buf, "k.k0 = x.k%zu.lo; k.k1 = x.k%zu.hi;", s, s);
          else snprintf(buf, sizeof buf, "%s |= (uint64_t)x.k%zu << %d;", word.c_str(), s, shift);
          break;
        case K_BOOL: snprintf(buf, sizeof buf, "%s |= (uint64_t)(x.k%zu ? 1 : 0) << %d;", word.c_str(), s, shift); break;
        case K_DICT32:
          snprintf(buf, sizeof buf, "if (x.k%zu < 0) k.meta |= %uu; else %s |= (uint64_t)(uint32_t)x.k%zu << %d;", s, 1u << ki,
                   word.c_str(), s, shift);
          break;
        default:
          snprintf(buf, sizeof buf, "too_long |= x.l%zu > %d; str_field<%d, %d>(a.col[%zu], x.b%zu, x.l%zu, x.w%zu, k.k0, k.k1);",
                   s, width, off, width, s, s, s, s);
          break;
      }
      body = buf;
      if (sl.nullable) o.fmt("    if (!x.v%zu) k.meta |= %uu; else { %s }", s, 1u << ki, body.c_str());
      else o.fmt("    { %s }", body.c_str());
    }
    o.line("    if (too_long) atomicOr(&a.t.flags[1], 1);");
    o.line("    return k;");
    o.line("  }");
  }

  // expression text of one atom in class `cls`; *ok_terms collects the validity terms
  int atom(const HAtom& h, __int128 mul, int cls, std::string* expr, std::vector<std::string>* ok_terms, std::string* err) {
    if (!h.is_col) {
      if (cls == CLS_F64) { *expr = f64_lit(h.flit); return BGPU_OK; }
      __int128 v = h.lit;
      if (cls == CLS_I128 && __builtin_mul_overflow(h.lit, mul, &v)) {
        *err = "Overflow happened on rescaling a decimal literal";
        return BGPU_EOVERFLOW;
      }
      *expr = cls == CLS_I128 ? val_lit(v) : hex64((uint64_t)v);
      return BGPU_OK;
    }
    const int s = slot_of(h.col);
    const JitSlot& sl = sh.slots[(size_t)s];
    if (sl.nullable) ok_terms->push_back("x.v" + std::to_string(s));
    if (cls == CLS_I128) {
      std::string e = as_i128(s, h.vk);
      if (mul != 1) {
        if (mul > (__int128)UINT64_MAX) { *err = "decimal rescale factor beyond 10^19"; return BGPU_EUNSUPPORTED; }
        e = "i128_mul_checked(" + e + ", Val{" + hex64((uint64_t)mul) + ", 0}, &o)";
      }
      *expr = e;
    } else if (cls == CLS_F64) {
      *expr = as_double(s, h.vk);
    } else {
      *expr = as_u64(s, h.vk);
    }
    return BGPU_OK;
  }

  int accumulate(std::string* err) {
    o.line("  template <class Sink> static __device__ __forceinline__ uint32_t accumulate(const JitArgs& a, const Raw& x, Sink& sink, bool* ovf) {");
    o.line("    uint32_t seen = 0;");
    for (size_t vi = 0; vi < g->values.size(); vi++) {
      const HValue& h = g->values[vi];
      const bool from_prev = g->value_from_prev[vi];
      const char* ty = h.cls == CLS_F64 ? "double" : h.cls == CLS_I128 ? "Val" : "uint64_t";
      std::vector<std::string> ok_terms;
      if (from_prev) ok_terms.push_back("ok" + std::to_string(vi - 1));
      std::vector<std::string> lines;
      if (h.cls != CLS_NONE) {
        std::string prod;
        for (size_t fi = from_prev ? h.f.size() - 1 : 0; fi < h.f.size(); fi++) {
          const HFactor& f = h.f[fi];
          std::string ex, ey, fe;
          int rc = atom(f.x, f.x_mul, h.cls, &ex, &ok_terms, err);
          if (rc) return rc;
          if (f.op) {
            rc = atom(f.y, f.y_mul, h.cls, &ey, &ok_terms, err);
            if (rc) return rc;
            if (h.cls == CLS_I128) fe = std::string(f.op == 1 ? "i128_add_checked(" : "i128_sub_checked(") + ex + ", " + ey + ", &o)";
            else fe = "(" + ex + (f.op == 1 ? " + " : " - ") + ey + ")";
          } else {
            fe = ex;
          }
          lines.push_back("      const " + std::string(ty) + " f" + std::to_string(fi) + " = " + fe + ";");
        }
        prod = from_prev ? "v" + std::to_string(vi - 1) : "f0";
        for (size_t fi = from_prev ? h.f.size() - 1 : 1; fi < h.f.size(); fi++) {
          const std::string f = "f" + std::to_string(fi);
          prod = h.cls == CLS_I128 ? "i128_mul_checked(" + prod + ", " + f + ", &o)" : "(" + prod + " * " + f + ")";
        }
        lines.push_back("      v" + std::to_string(vi) + " = " + prod + ";");
      }
      std::string ok = "true";
      for (const std::string& t : ok_terms) ok += " && " + t;
      o.fmt("    const bool ok%zu = %s;", vi, ok.c_str());
      if (h.cls != CLS_NONE) o.fmt("    %s v%zu = %s;", ty, vi, h.cls == CLS_I128 ? "Val{0, 0}" : "0");
      o.fmt("    if (ok%zu) {", vi);
      if (h.cls == CLS_I128) o.line("      bool o = false;");
      for (const std::string& l : lines) o.line(l);
      if (h.cls == CLS_I128) o.line("      if (o) *ovf = true;");
      std::string as_val = h.cls == CLS_NONE ? "Val{1, 0}"
                           : h.cls == CLS_I128 ? "v" + std::to_string(vi)
                           : h.cls == CLS_F64 ? "Val{f64_bits(v" + std::to_string(vi) + "), 0}"
                                              : "Val{v" + std::to_string(vi) + ", 0}";
      for (size_t ai = 0; ai < g->phys.size(); ai++) {
        if (g->phys[ai].value != (int32_t)vi) continue;
        const int alias = sh.acc_alias[ai];
        const int word0 = alias >= 0 ? sh.acc_word0[(size_t)alias] : sh.acc_word0[ai];
        if (alias >= 0) o.fmt("      if (!Sink::kSkipAliased) sink.template add<%d, %d, %zu, %d>(a, %s);", g->phys[ai].op, word0, ai, g->phys[ai].out_off, as_val.c_str());
        else o.fmt("      sink.template add<%d, %d, %zu, %d>(a, %s);", g->phys[ai].op, word0, ai, g->phys[ai].out_off, as_val.c_str());
        o.fmt("      seen |= %uu;", 1u << ai);
      }
      o.line("    }");
    }
    o.line("    return HAS_SEEN ? seen : 0u;");
    o.line("  }");
    return BGPU_OK;
  }

  void for_each_acc() {
    o.line("  template <class F> static __device__ __forceinline__ void for_each_acc(F& f) {");
    for (size_t ai = 0; ai < g->phys.size(); ai++) {
      const int alias = sh.acc_alias[ai];
      const int word0 = alias >= 0 ? sh.acc_word0[(size_t)alias] : sh.acc_word0[ai];
      o.fmt("    f.template acc<%d, %d, %d, %zu, %s, %d>();", g->phys[ai].op, word0, g->phys[ai].words, ai, alias >= 0 ? "true" : "false", g->phys[ai].out_off);
    }
    o.line("  }");
  }
};

}  // namespace

int agg_generate_source(const bgpu_agg* g, const JitShape& sh, std::string* out) {
  Gen gen{g, sh, {}};
  Out& o = gen.o;
  o.s.reserve(8192);
  o.line("namespace bgpu {");
  o.line("struct Plan {");
  o.fmt("  static constexpr int T = %d, R = %d, W = %d, KEY_WORDS = %d, MIN_CTAS = %d;", sh.T, sh.R, sh.W, g->key_words, sh.min_ctas);
  o.fmt("  static constexpr bool HAS_SEEN = %s;", sh.has_seen ? "true" : "false");
  o.fmt("  typedef %s Row;  // row index type of the loads", sh.row32 ? "uint32_t" : "int64_t");
  o.fmt("  static constexpr bool L2_PREFETCH = %s;", getenv("BGPU_L2_PREFETCH") ? "true" : "false");
  // one row per thread and tile: run the three-stage load pipeline (fused_body)
  const char* pipe = getenv("BGPU_FUSED_PIPE");
  o.fmt("  static constexpr bool PIPE = %s;", (pipe ? atoi(pipe) != 0 : sh.R == 1) ? "true" : "false");
  {  // thread-private layout of the fused kernel (PrivateSink): u64 slot and, for i128 sums, i32 high-word slot
    std::string p8 = "0", p4 = "0";
    int n8 = 0, n4 = 0;
    for (size_t ai = 0; ai < g->phys.size(); ai++) {
      if (sh.acc_alias[ai] >= 0) continue;
      p8 = "w == " + std::to_string(sh.acc_word0[ai]) + " ? " + std::to_string(n8++) + " : " + p8;
      if (g->phys[ai].words == 2) p4 = "w == " + std::to_string(sh.acc_word0[ai]) + " ? " + std::to_string(n4++) + " : " + p4;
    }
    o.fmt("  static constexpr int P8 = %d, P4 = %d, PRIV_WORDS = %d;", n8, n4, n8 + (n4 + 1) / 2);
    o.fmt("  static __device__ constexpr int p8_of(int w) { return %s; }", p8.c_str());
    o.fmt("  static __device__ constexpr int p4_of(int w) { return %s; }", p4.c_str());
  }
  gen.raw_struct();
  gen.load0();
  gen.prefetch();
  gen.load1();
  gen.load2();
  gen.filter();
  gen.key();
  std::string err;
  const int rc = gen.accumulate(&err);
  if (rc) return g->ctx->fail(rc, err);
  gen.for_each_acc();
  o.line("};");
  o.line("}  // namespace bgpu");
  o.line("using namespace bgpu;");
  if (sh.kernels & 1)
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T, Plan::MIN_CTAS) bgpu_fused(const __grid_constant__ JitArgs a) { fused_body<Plan, 0>(a); }");
  if (sh.kernels & 4)
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T, Plan::MIN_CTAS) bgpu_fused_shared(const __grid_constant__ JitArgs a) { fused_body<Plan, 1>(a); }");
  if (sh.kernels & 2) {
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T, Plan::MIN_CTAS) bgpu_intern(const __grid_constant__ JitArgs a) { intern_body<Plan>(a); }");
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T) bgpu_acc_shared(const __grid_constant__ JitArgs a) { accumulate_body<Plan, 1>(a); }");
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T, Plan::MIN_CTAS) bgpu_acc_global(const __grid_constant__ JitArgs a) { accumulate_body<Plan, 2>(a); }");
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T) bgpu_sketch(const __grid_constant__ JitArgs a) { sketch_body<Plan>(a); }");
    o.line("extern \"C\" __global__ void __launch_bounds__(Plan::T) bgpu_bucket(const __grid_constant__ JitArgs a) { bucket_body<Plan>(a); }");
  }
  *out = std::move(o.s);
  return BGPU_OK;
}

// Device-side building blocks of the grouped aggregate: key packing, the group table (global, plus a per-CTA shared
// memory cache in front of it), checked Decimal128 arithmetic, value/filter evaluation and accumulator adds.
#pragma once
#include "agg_plan.h"

namespace bgpu {

__device__ __forceinline__ bool bit_set(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // murmur3 fmix64: table hash (not the partition hash)
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

struct PackedKey { uint64_t k0, k1; uint32_t meta; };

__device__ __forceinline__ uint64_t key_hash(const PackedKey& k) {
  return mix64(k.k0 ^ mix64(k.k1 + 0x9e3779b97f4a7c15ULL * (k.meta + 1)));
}

__device__ __forceinline__ void put_bytes(PackedKey& k, int off, uint64_t v) {
  // place v at byte offset off of the 128-bit key (fields never straddle the 8-byte boundary except utf8, which is
  // put byte by byte)
  if (off < 8) k.k0 |= v << (8 * off); else k.k1 |= v << (8 * (off - 8));
}

__device__ __forceinline__ PackedKey pack_key(const KeyPlan& kp, int64_t r, int32_t* flags) {
  PackedKey k{0, 0, 0};
#pragma unroll 1
  for (int c = 0; c < kp.n; c++) {
    const KeyCol& kc = kp.cols[c];
    if (kc.validity && !bit_set(kc.validity, r)) { k.meta |= 1u << c; continue; }
    switch (kc.kind) {
      case K_DICT32: {  // keyed by the canonical dictionary index, i.e. by value (the reference keys on values)
        int32_t idx = ((const int32_t*)kc.values)[r];
        if (kc.remap) idx = kc.remap[idx];
        if (idx < 0) k.meta |= 1u << c;
        else put_bytes(k, kc.byte_off, (uint32_t)idx);
      } break;
      case K_FIXED:
        switch (kc.width) {
          case 8: put_bytes(k, kc.byte_off, ((const uint64_t*)kc.values)[r]); break;
          case 4: put_bytes(k, kc.byte_off, ((const uint32_t*)kc.values)[r]); break;
          case 2: put_bytes(k, kc.byte_off, ((const uint16_t*)kc.values)[r]); break;
          case 1: put_bytes(k, kc.byte_off, ((const uint8_t*)kc.values)[r]); break;
          default: {  // 16: the only key column
            const ulonglong2 v = ((const ulonglong2*)kc.values)[r];
            k.k0 = v.x; k.k1 = v.y;
          } break;
        }
        break;
      case K_BOOL: put_bytes(k, kc.byte_off, bit_set((const uint8_t*)kc.values, r) ? 1 : 0); break;
      default: {  // K_UTF8 / K_BINARY: length byte then up to `width` bytes
        const int32_t b = kc.offsets[r], len = kc.offsets[r + 1] - b;
        if (len > kc.width) { atomicOr(&flags[1], 1); break; }
        put_bytes(k, kc.byte_off, (uint64_t)len);
        const uint8_t* p = (const uint8_t*)kc.values + b;
        for (int i = 0; i < len; i++) put_bytes(k, kc.byte_off + 1 + i, p[i]);
      } break;
    }
  }
  return k;
}

// Find or insert in the global table; returns the dense group id (>= max_groups when the table overflowed: the
// pass is void and is re-run by the host after growing).
constexpr uint32_t kGidOverflow = 0xfffffff0u;  // "no room": always >= max_groups

__device__ __noinline__ uint32_t intern(const GroupTable& t, const PackedKey& k, int words) {
  const uint64_t h = key_hash(k);
  const unsigned long long tag = (h >> 32) << 32;
  uint32_t i = (uint32_t)h & t.mask;
  for (uint32_t probes = 0; probes <= t.mask; probes++) {
    unsigned long long e = *(volatile unsigned long long*)&t.entries[i];
    if (e == 0) {
      if (*(volatile int32_t*)&t.flags[0]) return kGidOverflow;  // the pass is void already: do not fill the table
      e = atomicCAS(&t.entries[i], 0ULL, tag | 1ULL);
      if (e == 0) {  // we own the slot: allocate the dense id, publish the key, then the entry
        const uint32_t gid = atomicAdd(t.n_groups, 1u);
        if (gid < t.max_groups) {
          t.key0[gid] = k.k0;
          if (words > 1) t.key1[gid] = k.k1;
          t.kmeta[gid] = k.meta;
        } else {
          atomicOr(&t.flags[0], 1);
        }
        __threadfence();
        *(volatile unsigned long long*)&t.entries[i] = tag | (unsigned long long)(gid + 2u);
        return gid;
      }
    }
    if ((e & 0xffffffff00000000ULL) == tag) {
      while ((uint32_t)e == 1u) e = *(volatile unsigned long long*)&t.entries[i];  // inserter is publishing
      const uint32_t gid = (uint32_t)e - 2u;
      if (gid >= t.max_groups) return gid;
      // dense key arrays share cache lines between groups: read through L2 (L1 may hold a stale line)
      if (__ldcg(&t.key0[gid]) == k.k0 && (words == 1 || __ldcg(&t.key1[gid]) == k.k1) && __ldcg(&t.kmeta[gid]) == k.meta)
        return gid;
    }
    i = (i + 1) & t.mask;
  }
  atomicOr(&t.flags[0], 1);  // table full
  return kGidOverflow;
}

// Per-CTA cache of (key -> gid) in shared memory in front of the global table: low-cardinality inputs resolve
// every row from shared memory after the first few rows of the CTA.
struct KeyCache {
  uint32_t* state;  // 0 empty, 1 busy, gid + 2
  uint64_t* k0;
  uint64_t* k1;
  uint32_t* meta;
  uint32_t mask;  // slots - 1
};
__device__ __forceinline__ size_t key_cache_bytes(uint32_t slots) { return (size_t)slots * (4 + 8 + 8 + 4); }

__device__ __forceinline__ uint32_t cached_intern(const KeyCache& c, const GroupTable& t, const PackedKey& k, int words) {
  uint32_t i = (uint32_t)(key_hash(k) >> 7) & c.mask;
#pragma unroll 1
  for (uint32_t probe = 0; probe <= c.mask; probe++) {
    const uint32_t s = *(volatile uint32_t*)&c.state[i];
    if (s >= 2u) {
      if (*(volatile uint64_t*)&c.k0[i] == k.k0 && (words == 1 || *(volatile uint64_t*)&c.k1[i] == k.k1) &&
          *(volatile uint32_t*)&c.meta[i] == k.meta)
        return s - 2u;
      i = (i + 1) & c.mask;
      continue;
    }
    const uint32_t gid = intern(t, k, words);
    if (s == 0u && probe * 2 < c.mask && atomicCAS(&c.state[i], 0u, 1u) == 0u) {
      c.k0[i] = k.k0;
      c.k1[i] = k.k1;
      c.meta[i] = k.meta;
      __threadfence_block();
      *(volatile uint32_t*)&c.state[i] = gid + 2u;
    }
    return gid;
  }
  return intern(t, k, words);
}

// ---- values ---------------------------------------------------------------------------------------------------
struct Val { uint64_t lo, hi; };  // i128, or i64 in lo, or the bits of a double in lo

__device__ __forceinline__ Val load_raw(const void* col, int vk, int64_t r) {
  Val v{0, 0};
  switch (vk) {
    case V_I128: { const ulonglong2 x = ((const ulonglong2*)col)[r]; v.lo = x.x; v.hi = x.y; } break;
    case V_I64: case V_U64: case V_F64: v.lo = ((const uint64_t*)col)[r]; break;
    case V_I32: v.lo = (uint64_t)(int64_t)((const int32_t*)col)[r]; break;
    case V_U32: v.lo = ((const uint32_t*)col)[r]; break;
    case V_F32: v.lo = (uint64_t)__double_as_longlong((double)((const float*)col)[r]); break;
    case V_I16: v.lo = (uint64_t)(int64_t)((const int16_t*)col)[r]; break;
    case V_U16: v.lo = ((const uint16_t*)col)[r]; break;
    case V_I8: v.lo = (uint64_t)(int64_t)((const int8_t*)col)[r]; break;
    case V_U8: v.lo = ((const uint8_t*)col)[r]; break;
    default: break;
  }
  return v;
}
__device__ __forceinline__ bool vk_signed_int(int vk) { return vk == V_I8 || vk == V_I16 || vk == V_I32 || vk == V_I64; }
__device__ __forceinline__ bool vk_unsigned_int(int vk) { return vk == V_U8 || vk == V_U16 || vk == V_U32 || vk == V_U64; }

// i128 helpers on (lo, hi) pairs
__device__ __forceinline__ Val i128_neg(Val a) {
  Val r;
  r.lo = ~a.lo + 1;
  r.hi = ~a.hi + (r.lo == 0);
  return r;
}
__device__ __forceinline__ bool i128_is_neg(Val a) { return (int64_t)a.hi < 0; }
// checked add: returns false on signed overflow
__device__ __forceinline__ bool i128_add_checked(Val a, Val b, Val* out) {
  Val r;
  r.lo = a.lo + b.lo;
  r.hi = a.hi + b.hi + (r.lo < a.lo);
  *out = r;
  return !((~(a.hi ^ b.hi) & (a.hi ^ r.hi)) >> 63);
}
__device__ __forceinline__ bool i128_sub_checked(Val a, Val b, Val* out) {
  Val r;
  r.lo = a.lo - b.lo;
  r.hi = a.hi - b.hi - (a.lo < b.lo);
  *out = r;
  return !(((a.hi ^ b.hi) & (a.hi ^ r.hi)) >> 63);
}
// checked signed multiply (i128::mul_checked): magnitudes as u128, product must fit 127 bits (or be exactly MIN)
__device__ __forceinline__ bool i128_mul_checked(Val a, Val b, Val* out) {
  const bool neg = i128_is_neg(a) != i128_is_neg(b);
  const Val ua = i128_is_neg(a) ? i128_neg(a) : a, ub = i128_is_neg(b) ? i128_neg(b) : b;
  bool ok = (ua.hi == 0) || (ub.hi == 0);
  // (ua.hi*2^64 + ua.lo) * (ub.hi*2^64 + ub.lo), with at most one non-zero high word
  const uint64_t p0_lo = ua.lo * ub.lo, p0_hi = __umul64hi(ua.lo, ub.lo);
  const uint64_t cross_a = ua.hi * ub.lo, cross_a_hi = __umul64hi(ua.hi, ub.lo);
  const uint64_t cross_b = ua.lo * ub.hi, cross_b_hi = __umul64hi(ua.lo, ub.hi);
  ok = ok && cross_a_hi == 0 && cross_b_hi == 0;
  uint64_t hi = p0_hi + cross_a;
  ok = ok && hi >= p0_hi;
  const uint64_t hi2 = hi + cross_b;
  ok = ok && hi2 >= hi;
  Val r{p0_lo, hi2};
  if (neg) {
    ok = ok && (hi2 < 0x8000000000000000ULL || (hi2 == 0x8000000000000000ULL && p0_lo == 0));
    r = i128_neg(r);
  } else {
    ok = ok && hi2 < 0x8000000000000000ULL;
  }
  *out = r;
  return ok;
}
__device__ __forceinline__ bool i128_lt(Val a, Val b) {
  return (int64_t)a.hi < (int64_t)b.hi || (a.hi == b.hi && a.lo < b.lo);
}

// one atom of class `cls`: ok=false when null. Integers are sign-/zero-extended to i128 for CLS_I128.
__device__ __forceinline__ Val eval_atom(const Atom& a, int cls, int64_t r, bool* ok, bool* ovf) {
  if (!a.col) return Val{a.lit_lo, a.lit_hi};
  if (a.validity && !bit_set(a.validity, r)) { *ok = false; return Val{0, 0}; }
  Val v = load_raw(a.col, a.vk, r);
  if (cls == CLS_I128) {
    if (a.vk != V_I128) v.hi = vk_signed_int(a.vk) ? (uint64_t)((int64_t)v.lo >> 63) : 0;
    if (a.mul != 1) {
      Val m;
      if (!i128_mul_checked(v, Val{a.mul, 0}, &m)) *ovf = true;
      v = m;
    }
  } else if (cls == CLS_F64) {
    if (vk_signed_int(a.vk)) v.lo = (uint64_t)__double_as_longlong((double)(int64_t)v.lo);
    else if (vk_unsigned_int(a.vk)) v.lo = (uint64_t)__double_as_longlong((double)v.lo);
  }
  return v;
}

__device__ __forceinline__ Val eval_factor(const Factor& f, int cls, int64_t r, bool* ok, bool* ovf) {
  Val x = eval_atom(f.x, cls, r, ok, ovf);
  if (f.op == 0) return x;
  const Val y = eval_atom(f.y, cls, r, ok, ovf);
  if (cls == CLS_I128) {
    Val o;
    const bool fine = f.op == 1 ? i128_add_checked(x, y, &o) : i128_sub_checked(x, y, &o);
    if (!fine) *ovf = true;
    return o;
  }
  if (cls == CLS_F64) {
    const double a = __longlong_as_double((long long)x.lo), b = __longlong_as_double((long long)y.lo);
    return Val{(uint64_t)__double_as_longlong(f.op == 1 ? a + b : a - b), 0};
  }
  return Val{f.op == 1 ? x.lo + y.lo : x.lo - y.lo, 0};  // Int64: wrapping
}

__device__ __forceinline__ Val mul_class(Val a, Val b, int cls, bool* ovf) {
  if (cls == CLS_I128) {
    Val o;
    if (!i128_mul_checked(a, b, &o)) *ovf = true;
    return o;
  }
  if (cls == CLS_F64)
    return Val{(uint64_t)__double_as_longlong(__longlong_as_double((long long)a.lo) * __longlong_as_double((long long)b.lo)), 0};
  return Val{a.lo * b.lo, 0};
}

// value of expression e at row r; `prev`/`prev_ok` = the previous expression's result for this row
__device__ __forceinline__ Val eval_value(const ValueExpr& e, int64_t r, Val prev, bool prev_ok, bool* ok, bool* ovf) {
  if (e.cls == CLS_NONE) return Val{1, 0};
  Val v;
  int start;
  bool o = true;
  if (e.from_prev) {  // only the last factor is new
    v = prev;
    o = prev_ok;
    start = e.n_factors - 1;
  } else {
    v = eval_factor(e.f[0], e.cls, r, &o, ovf);
    start = 1;
  }
#pragma unroll 1
  for (int i = start; i < e.n_factors; i++) {
    const Val f = eval_factor(e.f[i], e.cls, r, &o, ovf);
    v = mul_class(v, f, e.cls, ovf);
  }
  *ok = o;
  return v;
}

__device__ __forceinline__ bool eval_filter(const FilterPlan& fp, int64_t r) {
#pragma unroll 1
  for (int i = 0; i < fp.n; i++) {
    const FilterTerm& t = fp.t[i];
    if (t.validity && !bit_set(t.validity, r)) return false;
    Val v = load_raw(t.col, t.vk, r);
    int c;  // -1, 0, 1
    if (t.vk == V_F64 || t.vk == V_F32) {
      const double a = __longlong_as_double((long long)v.lo), b = __longlong_as_double((long long)t.lit_lo);
      if (a != a) return false;  // NaN compares false (arrow's total order differs; NaN filters are off the path)
      c = a < b ? -1 : (a > b ? 1 : 0);
    } else {
      if (t.vk != V_I128) v.hi = vk_signed_int(t.vk) ? (uint64_t)((int64_t)v.lo >> 63) : 0;
      const Val l{t.lit_lo, t.lit_hi};
      c = i128_lt(v, l) ? -1 : (i128_lt(l, v) ? 1 : 0);
    }
    bool pass;
    switch (t.cmp) {
      case BGPU_X_LT: pass = c < 0; break;
      case BGPU_X_LE: pass = c <= 0; break;
      case BGPU_X_GT: pass = c > 0; break;
      case BGPU_X_GE: pass = c >= 0; break;
      case BGPU_X_EQ: pass = c == 0; break;
      default: pass = c != 0; break;
    }
    if (!pass) return false;
  }
  return true;
}

// ---- accumulator adds ---------------------------------------------------------------------------------------------
// plain (non-atomic) add into a thread-private slot; `stride` = distance between the words of one accumulator
__device__ __forceinline__ void add_private(uint64_t* p, int stride, int op, const Val& v) {
  if (op == OP_ADD_F64) {
    *(double*)p += __longlong_as_double((long long)v.lo);
  } else if (op == OP_ADD_I128) {
    const uint64_t old = p[0], lo = old + v.lo;
    p[0] = lo;
    p[stride] += v.hi + (lo < old);
  } else if (op == OP_COUNT) {
    p[0] += 1;
  } else {
    p[0] += v.lo;
  }
}
// atomic add (shared or global memory); i128 = two u64 adds, the carry out of the low word goes to the high word
__device__ __forceinline__ void add_atomic(uint64_t* p, int op, const Val& v) {
  if (op == OP_ADD_F64) {
    atomicAdd((double*)p, __longlong_as_double((long long)v.lo));
  } else if (op == OP_ADD_I128) {
    const uint64_t old = atomicAdd((unsigned long long*)p, (unsigned long long)v.lo);
    const uint64_t hi = v.hi + ((old + v.lo) < old);
    if (hi) atomicAdd((unsigned long long*)(p + 1), (unsigned long long)hi);
  } else if (op == OP_COUNT) {
    atomicAdd((unsigned long long*)p, 1ULL);
  } else {
    atomicAdd((unsigned long long*)p, (unsigned long long)v.lo);
  }
}

}  // namespace bgpu

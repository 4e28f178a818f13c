// Device code of the grouped aggregate (K4 Partial / K5 Final): the kernel templates and their building blocks.
// This file is compiled at run time by NVRTC for sm_100a (jit.cu) with a plan-specific `struct Plan` appended by
// agg_codegen.cu; it is embedded into libbgpu.so as a string (Makefile). It therefore uses no host headers.
//
//   fused_body<Plan>       ONE pass over the batch for low-cardinality inputs (TPC-H Q1, db-benchmark K=100):
//                          phase 1 issues every column load of R rows per thread back to back (memory-level
//                          parallelism), phase 2 filters, packs the key, resolves the group id from the CTA's
//                          shared-memory key cache (global table behind it), evaluates the aggregate arguments and adds
//                          into THREAD-PRIVATE accumulators in shared memory (no atomics, deterministic). Each CTA
//                          reduces its private copies into a scratch slab; commit_kernel (aggregate_run.cu) folds the
//                          slabs into the dense accumulator arrays once the host has seen that no CTA aborted.
//   intern_body<Plan> +    the general two-pass path: pass 1 assigns group ids (idempotent, so a table overflow is
//   accumulate_body<Plan>  handled by growing and re-running), pass 2 adds with replicated shared-memory atomics
//                          (TIER 1) or global atomics (TIER 2).
typedef unsigned char uint8_t;
typedef signed char int8_t;
typedef unsigned short uint16_t;
typedef short int16_t;
typedef unsigned int uint32_t;
typedef int int32_t;
typedef unsigned long long uint64_t;
typedef long long int64_t;
typedef unsigned long size_t;

namespace bgpu {

enum AccOp : int32_t { OP_ADD_I64 = 0, OP_ADD_I128 = 1, OP_ADD_F64 = 2, OP_COUNT = 3 };
constexpr uint32_t kNoGroup = 0xffffffffu;      // row dropped by the filter
constexpr uint32_t kGidOverflow = 0xfffffff0u;  // "no room": always >= max_groups
constexpr int kJitSlots = 24;
constexpr int kJitAcc = 24;

// The global group table (host mirror: agg_plan.h, same layout): open addressing with the packed key stored INSIDE
// the slot, so a probe is one 16-byte (one key word) or 32-byte (two key words) access — one DRAM sector.
//   slot = [k0][k1 if KEY_WORDS == 2][sm][pad if KEY_WORDS == 2],  sm = (null bits << 32) | state,
//   state: 0 empty, 1 busy (being published), >= 2: dense group id + 2.
// Accumulators: one record of acc_stride u64 per group (array of structs).
struct GroupTable {
  uint64_t* slots;
  uint32_t mask;        // capacity - 1 (capacity = 2 * max_groups, power of two)
  uint32_t max_groups;
  uint32_t* n_groups;   // device counter
  int32_t* flags;       // [0] capacity overflow, [1] string key too long, [2] arithmetic overflow,
                        // [3] fused pass abort (a group id beyond the per-CTA capacity)
  uint64_t* acc;        // max_groups * acc_stride u64
  uint32_t acc_stride;
  uint32_t slot_words;  // 2 or 4
};

// The single by-value kernel argument; the host mirror is agg_plan.h's JitArgs (same layout).
struct JitArgs {
  const void* col[kJitSlots];        // slot s: fixed values / bool bits / utf8 bytes / dictionary keys
  const int32_t* aux[kJitSlots];     // slot s: utf8 offsets, or the dictionary remap (nullptr = identity)
  const uint8_t* valid[kJitSlots];   // slot s: validity bitmap (only read when the plan marks the slot nullable)
  GroupTable t;
  uint32_t* seen;                    // per-group bitmask of accumulators that saw a non-null input
  uint64_t* scratch;                 // fused path: per-CTA slabs
  uint32_t* gids;                    // two-pass path: group id per row
  int64_t n_rows;
  uint32_t g_cap;                    // fused: private group capacity (ids >= g_cap abort the pass)
  uint32_t cache_slots;              // fused: key cache slots (power of two)
  uint32_t n_groups;                 // accumulate: groups so far
  uint32_t replicas;                 // accumulate TIER 1: shared-memory copies per CTA
  uint16_t* bucket_ids;              // radix pre-pass: table bucket per row
  uint32_t* bucket_hist;             // radix pre-pass: hist[bucket][chunk]
  uint32_t bucket_shift;             // bucket = (table slot index) >> bucket_shift
  uint32_t bucket_chunk;             // rows per CTA chunk (the scatter's chunk layout)
};

__device__ __forceinline__ bool bit_set(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // murmur3 fmix64: table hash (not the partition hash)
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

struct PackedKey { uint64_t k0, k1; uint32_t meta; };

template <int WORDS>
__device__ __forceinline__ uint64_t key_hash(const PackedKey& k) {
  // must equal agg_device.cuh's key_hash (rehash_kernel rebuilds the table with it); k1 == 0 when WORDS == 1
  return mix64(k.k0 ^ mix64((WORDS > 1 ? k.k1 : 0ULL) + 0x9e3779b97f4a7c15ULL * (k.meta + 1)));
}

// Find or insert in the global table; returns the dense group id (>= max_groups when the table overflowed: the
// pass is void and is re-run by the host after growing).
template <int WORDS>
__device__ __noinline__ uint32_t intern_slow(const GroupTable& t, uint64_t k0, uint64_t k1, uint32_t meta) {
  const PackedKey k{k0, k1, meta};  // by value: a reference would force the caller's key into local memory
  constexpr int SW = WORDS == 1 ? 2 : 4, SMW = WORDS == 1 ? 1 : 2;  // slot words, index of the state|meta word
  uint32_t i = (uint32_t)key_hash<WORDS>(k) & t.mask;
  for (uint32_t probes = 0; probes <= t.mask; probes++) {
    uint64_t* slot = t.slots + (size_t)i * SW;
    // one 16-byte access brings key word 0 and state|meta (WORDS == 1) or both key words (WORDS == 2); L2-coherent
    const ulonglong2 a = __ldcg((const ulonglong2*)slot);
    uint64_t sm = WORDS == 1 ? a.y : __ldcg(slot + SMW);
    if ((uint32_t)sm == 0u) {
      if (*(volatile int32_t*)&t.flags[0]) return kGidOverflow;  // the pass is void already: do not fill the table
      const unsigned long long prev = atomicCAS((unsigned long long*)(slot + SMW), 0ULL, 1ULL);
      if (prev == 0ULL) {  // we own the slot: write the key, allocate the dense id, then publish state|meta
        slot[0] = k.k0;
        if (WORDS > 1) slot[1] = k.k1;
        const uint32_t gid = atomicAdd(t.n_groups, 1u);
        if (gid >= t.max_groups) atomicOr(&t.flags[0], 1);
        __threadfence();
        *(volatile uint64_t*)(slot + SMW) = ((uint64_t)k.meta << 32) | (uint64_t)(gid + 2u);
        return gid;
      }
      sm = prev;
    }
    while ((uint32_t)sm == 1u) sm = *(volatile uint64_t*)(slot + SMW);  // inserter is publishing
    // published slots never change again. The key words above may predate the publication we just observed (or the
    // 16-byte access may have been split): on a mismatch re-read them once before moving on.
    if ((uint32_t)(sm >> 32) == k.meta) {
      bool same = a.x == k.k0 && (WORDS == 1 || a.y == k.k1);
      if (!same) same = *(volatile uint64_t*)slot == k.k0 && (WORDS == 1 || *(volatile uint64_t*)(slot + 1) == k.k1);
      if (same) return (uint32_t)sm - 2u;
    }
    i = (i + 1) & t.mask;
  }
  atomicOr(&t.flags[0], 1);  // table full
  return kGidOverflow;
}
template <int WORDS>
__device__ __forceinline__ uint32_t intern(const GroupTable& t, const PackedKey& k) {
  return intern_slow<WORDS>(t, k.k0, WORDS > 1 ? k.k1 : 0ULL, k.meta);
}

// Warp-synchronous find-or-insert of R keys per lane (intern pass of the two-pass path; all 32 lanes call it).
// Two things the one-key-at-a-time routine above cannot do at 10^7..10^8 groups:
//   * the slot loads of all R rows of a lane are in flight together (a probe is a DRAM round trip), and
//   * dense ids are allocated with ONE atomicAdd per warp and round for all lanes that claimed a slot in it - a
//     single-address atomic per new group serialises in L2 (measured: 25 M inserts = 30 ms).
// Rounds: load the slot of every unresolved row; match -> done, other key -> next slot, empty -> claim it (CAS), busy
// -> look again next round. Claimed slots get their ids, keys and state|meta published before the next round, and
// nobody ever waits inside a round, so lanes of one warp holding the same new key cannot deadlock.
template <int WORDS, int R>
__device__ __forceinline__ void warp_intern(const GroupTable& t, const PackedKey (&k)[R], const bool (&active)[R],
                                            uint32_t (&gid)[R]) {
  constexpr int SW = WORDS == 1 ? 2 : 4, SMW = WORDS == 1 ? 1 : 2;
  const uint32_t lane = threadIdx.x & 31, lt_mask = (1u << lane) - 1;
  uint32_t idx[R], probes[R];
  bool pend[R];
  bool any = false;
  // read once per call, not per empty slot: every SM polling one address is an L2 hot spot (25 M reads = 10+ ms)
  const bool overflowed = *(volatile int32_t*)&t.flags[0] != 0;
#pragma unroll
  for (int u = 0; u < R; u++) {
    idx[u] = (uint32_t)key_hash<WORDS>(k[u]) & t.mask;
    probes[u] = 0;
    pend[u] = active[u];
    gid[u] = kNoGroup;
    any |= pend[u];
  }
  while (__any_sync(0xffffffffu, any)) {
    ulonglong2 sa[R];
    uint64_t sm[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      if (!pend[u]) continue;
      const uint64_t* slot = t.slots + (size_t)idx[u] * SW;
      sa[u] = __ldcg((const ulonglong2*)slot);
      sm[u] = WORDS == 1 ? sa[u].y : __ldcg(slot + SMW);
    }
    any = false;
    bool won[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      won[u] = false;
      if (!pend[u]) continue;
      uint64_t* slot = t.slots + (size_t)idx[u] * SW;
      const uint32_t state = (uint32_t)sm[u];
      if (state >= 2u) {
        bool same = (uint32_t)(sm[u] >> 32) == k[u].meta;
        if (same) {
          same = sa[u].x == k[u].k0 && (WORDS == 1 || sa[u].y == k[u].k1);
          // the key words may predate the publication just observed: re-read them once before moving on
          if (!same) same = *(volatile uint64_t*)slot == k[u].k0 && (WORDS == 1 || *(volatile uint64_t*)(slot + 1) == k[u].k1);
        }
        if (same) {
          gid[u] = state - 2u;
          pend[u] = false;
        } else if (++probes[u] > t.mask) {
          atomicOr(&t.flags[0], 1);  // table full
          gid[u] = kGidOverflow;
          pend[u] = false;
        } else {
          idx[u] = (idx[u] + 1) & t.mask;
        }
      } else if (state == 0u) {
        if (overflowed) {  // the pass is void already: do not fill the table
          gid[u] = kGidOverflow;
          pend[u] = false;
        } else {
          won[u] = atomicCAS((unsigned long long*)(slot + SMW), 0ULL, 1ULL) == 0ULL;
        }
      }  // state == 1: an inserter is publishing; look again next round
    }
    // one id allocation per warp and round for every slot claimed in it
    uint32_t winners[R], before[R], total = 0;
#pragma unroll
    for (int u = 0; u < R; u++) {
      winners[u] = __ballot_sync(0xffffffffu, won[u]);
      before[u] = total;
      total += __popc(winners[u]);
    }
    if (total) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(t.n_groups, total);
      base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
      for (int u = 0; u < R; u++) {
        if (!won[u]) continue;
        uint64_t* slot = t.slots + (size_t)idx[u] * SW;
        gid[u] = base + before[u] + __popc(winners[u] & lt_mask);
        slot[0] = k[u].k0;
        if (WORDS > 1) slot[1] = k[u].k1;
        if (gid[u] >= t.max_groups) atomicOr(&t.flags[0], 1);
      }
      __threadfence();
#pragma unroll
      for (int u = 0; u < R; u++) {
        if (!won[u]) continue;
        uint64_t* slot = t.slots + (size_t)idx[u] * SW;
        *(volatile uint64_t*)(slot + SMW) = ((uint64_t)k[u].meta << 32) | (uint64_t)(gid[u] + 2u);
        pend[u] = false;
      }
    }
#pragma unroll
    for (int u = 0; u < R; u++) any |= pend[u];
  }
}

// Per-CTA cache of (key -> gid) in shared memory in front of the global table: low-cardinality inputs resolve
// every row from shared memory after the first few rows of the CTA.
struct KeyCache {
  // entries of 16 (one key word) or 24 bytes: k0, [k1], state (0 empty, 1 busy, gid + 2), meta
  unsigned char* base;
  uint32_t mask;  // slots - 1
};

template <int WORDS>
__device__ __forceinline__ uint32_t cached_intern(const KeyCache& c, const GroupTable& t, const PackedKey& k) {
  constexpr uint32_t ENTRY = WORDS == 1 ? 16 : 24, STATE = ENTRY - 8;
  // cheap slot hash in 32-bit multiplies: the cache holds a handful of keys, the strong hash is only needed behind it
  uint32_t h = (uint32_t)k.k0 * 0x9e3779b1u ^ (uint32_t)(k.k0 >> 32) * 0x85ebca6bu ^ k.meta;
  if (WORDS > 1) h ^= (uint32_t)k.k1 * 0xc2b2ae35u ^ (uint32_t)(k.k1 >> 32) * 0x27d4eb2fu;
  uint32_t i = ((h * 0x165667b1u) >> 15) & c.mask;
#pragma unroll 1
  for (uint32_t probe = 0; probe <= c.mask; probe++) {
    unsigned char* e = c.base + i * ENTRY;
    const uint32_t s = *(volatile uint32_t*)(e + STATE);
    if (s >= 2u) {
      if (*(volatile uint64_t*)e == k.k0 && (WORDS == 1 || *(volatile uint64_t*)(e + 8) == k.k1) &&
          *(volatile uint32_t*)(e + STATE + 4) == k.meta)
        return s - 2u;
      i = (i + 1) & c.mask;
      continue;
    }
    const uint32_t gid = intern<WORDS>(t, k);
    if (s == 0u && probe * 2 < c.mask && atomicCAS((uint32_t*)(e + STATE), 0u, 1u) == 0u) {
      *(uint64_t*)e = k.k0;
      if (WORDS > 1) *(uint64_t*)(e + 8) = k.k1;
      *(uint32_t*)(e + STATE + 4) = k.meta;
      __threadfence_block();
      *(volatile uint32_t*)(e + STATE) = gid + 2u;
    }
    return gid;
  }
  return intern<WORDS>(t, k);
}

// ---- values ---------------------------------------------------------------------------------------------------
struct Val { uint64_t lo, hi; };  // i128, or i64 in lo, or the bits of a double in lo

template <class I>
__device__ __forceinline__ Val ld_i128(const void* col, I r) {
  const ulonglong2 x = __ldg((const ulonglong2*)col + r);
  return Val{x.x, x.y};
}
__device__ __forceinline__ Val sext(int64_t v) { return Val{(uint64_t)v, (uint64_t)(v >> 63)}; }
__device__ __forceinline__ Val zext(uint64_t v) { return Val{v, 0}; }
__device__ __forceinline__ double as_f64(uint64_t b) { return __longlong_as_double((long long)b); }
__device__ __forceinline__ uint64_t f64_bits(double d) { return (uint64_t)__double_as_longlong(d); }

// 128-bit add / subtract as one carry chain (add.cc / addc): four 32-bit adds in SASS, no compare-and-select
__device__ __forceinline__ Val add128(Val a, Val b) {
  Val r;
  asm("add.cc.u64 %0, %2, %4;\n\taddc.u64 %1, %3, %5;" : "=l"(r.lo), "=l"(r.hi) : "l"(a.lo), "l"(a.hi), "l"(b.lo), "l"(b.hi));
  return r;
}
__device__ __forceinline__ Val sub128(Val a, Val b) {
  Val r;
  asm("sub.cc.u64 %0, %2, %4;\n\tsubc.u64 %1, %3, %5;" : "=l"(r.lo), "=l"(r.hi) : "l"(a.lo), "l"(a.hi), "l"(b.lo), "l"(b.hi));
  return r;
}
__device__ __forceinline__ Val i128_neg(Val a) { return sub128(Val{0, 0}, a); }
__device__ __forceinline__ bool i128_is_neg(Val a) { return (int64_t)a.hi < 0; }
__device__ __forceinline__ bool fits_i64(Val a) { return a.hi == (uint64_t)((int64_t)a.lo >> 63); }
// checked add/sub: *ovf is set on signed overflow (arrow-arith add_checked / sub_checked on i128)
__device__ __forceinline__ Val i128_add_checked(Val a, Val b, bool* ovf) {
  const Val r = add128(a, b);
  if ((~(a.hi ^ b.hi) & (a.hi ^ r.hi)) >> 63) *ovf = true;
  return r;
}
__device__ __forceinline__ Val i128_sub_checked(Val a, Val b, bool* ovf) {
  const Val r = sub128(a, b);
  if (((a.hi ^ b.hi) & (a.hi ^ r.hi)) >> 63) *ovf = true;
  return r;
}
// general checked signed multiply (i128::mul_checked): magnitudes as u128, product must fit 127 bits (or be MIN).
// Returns the overflow flag by value (a pointer argument would force the caller's flag into local memory).
struct ValOvf { Val v; uint32_t ovf; };
__device__ __noinline__ ValOvf i128_mul_slow(Val a, Val b) {
  const bool neg = i128_is_neg(a) != i128_is_neg(b);
  const Val ua = i128_is_neg(a) ? i128_neg(a) : a, ub = i128_is_neg(b) ? i128_neg(b) : b;
  bool ok = (ua.hi == 0) || (ub.hi == 0);
  const uint64_t p0_lo = ua.lo * ub.lo, p0_hi = __umul64hi(ua.lo, ub.lo);
  const uint64_t cross_a = ua.hi * ub.lo, cross_a_hi = __umul64hi(ua.hi, ub.lo);
  const uint64_t cross_b = ua.lo * ub.hi, cross_b_hi = __umul64hi(ua.lo, ub.hi);
  ok = ok && cross_a_hi == 0 && cross_b_hi == 0;
  const uint64_t hi = p0_hi + cross_a;
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
  return ValOvf{r, ok ? 0u : 1u};
}
// checked multiply with the common cases inline: both operands fit i32, or both fit i64 (one signed 64x64->128
// product, which cannot overflow i128)
__device__ __forceinline__ Val i128_mul_checked(Val a, Val b, bool* ovf) {
  {  // both operands fit i32 (prices x percentages): a single 32x32->64 multiply
    const uint32_t sa = (uint32_t)((int32_t)(uint32_t)a.lo >> 31), sb = (uint32_t)((int32_t)(uint32_t)b.lo >> 31);
    const uint32_t wide = (((uint32_t)(a.lo >> 32) ^ sa) | ((uint32_t)a.hi ^ sa) | ((uint32_t)(a.hi >> 32) ^ sa)) |
                          (((uint32_t)(b.lo >> 32) ^ sb) | ((uint32_t)b.hi ^ sb) | ((uint32_t)(b.hi >> 32) ^ sb));
    if (wide == 0) {
      const int64_t p = (int64_t)(int32_t)(uint32_t)a.lo * (int64_t)(int32_t)(uint32_t)b.lo;
      return Val{(uint64_t)p, (uint64_t)(p >> 63)};
    }
  }
  if (fits_i64(a) && fits_i64(b)) return Val{a.lo * b.lo, (uint64_t)__mul64hi((int64_t)a.lo, (int64_t)b.lo)};
  const ValOvf r = i128_mul_slow(a, b);
  if (r.ovf) *ovf = true;
  return r.v;
}
__device__ __forceinline__ bool i128_lt(Val a, Val b) {
  return (int64_t)a.hi < (int64_t)b.hi || (a.hi == b.hi && a.lo < b.lo);
}
// three-way compare of i128 values: -1, 0, 1
__device__ __forceinline__ int i128_cmp(Val a, Val b) { return i128_lt(a, b) ? -1 : (i128_lt(b, a) ? 1 : 0); }
__device__ __forceinline__ int f64_cmp(double a, double b) { return a < b ? -1 : (a > b ? 1 : 0); }

// ---- utf8 / binary group keys ------------------------------------------------------------------------------------------
// The key field of a string is [length byte][bytes...] placed at byte OFF of the 16-byte packed key. The bytes are
// fetched as the aligned 8-byte words that contain them (one LDG.64 for short strings instead of a dependent byte
// loop); words that hold no byte of the string are not read, so no access leaves the 8-byte-aligned extent of the
// string itself.
// Ask the L2 to fetch [p, p + bytes) from HBM (TMA-style bulk prefetch, one instruction per buffer and tile). The
// range is shrunk to 16-byte granules inside itself, so nothing outside the caller's rows is touched.
__device__ __forceinline__ void bulk_prefetch_l2(const char* p, int64_t bytes) {
  const uint64_t lo = ((uint64_t)p + 15) & ~15ULL, hi = ((uint64_t)p + (uint64_t)bytes) & ~15ULL;
  if (hi > lo) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(lo), "r"((uint32_t)(hi - lo)) : "memory");
}

struct StrWords { uint64_t w0, w1, w2; };
// stage 1 (load phase): just the loads, so that the words of every string of the tile are in flight together
template <int MAXLEN>
__device__ __forceinline__ StrWords str_load(const void* data, int32_t b, int32_t len) {
  const uint64_t addr = (uint64_t)data + (uint64_t)(int64_t)b;
  const uint64_t* w = (const uint64_t*)(addr & ~7ULL);
  const int end = (int)(addr & 7) + len;
  StrWords r;
  r.w0 = (len > 0 && len <= MAXLEN) ? __ldg(w) : 0ULL;
  r.w1 = (end > 8 && len <= MAXLEN) ? __ldg(w + 1) : 0ULL;
  r.w2 = (MAXLEN + 7 > 16 && end > 16 && len <= MAXLEN) ? __ldg(w + 2) : 0ULL;
  return r;
}
// stage 2: assemble the key field; strings longer than MAXLEN contribute nothing (the caller raises the flag)
template <int OFF, int MAXLEN>
__device__ __forceinline__ void str_field(const void* data, int32_t b, int32_t len, const StrWords& r, uint64_t& k0, uint64_t& k1) {
  if (len > MAXLEN) return;
  const int sh = (int)(((uint64_t)data + (uint64_t)(int64_t)b) & 7) * 8;
  uint64_t lo = sh ? (r.w0 >> sh) | (r.w1 << (64 - sh)) : r.w0;
  uint64_t hi = sh ? (r.w1 >> sh) | (r.w2 << (64 - sh)) : r.w1;
  if (len < 8) { lo &= (1ULL << (8 * len)) - 1; hi = 0; }
  else if (len < 16) hi &= (1ULL << (8 * (len - 8))) - 1;
  // F = len | bytes << 8 (len <= 15 bytes: fits 128 bits), then F << 8 * OFF
  const uint64_t f_lo = (uint64_t)len | (lo << 8), f_hi = (hi << 8) | (lo >> 56);
  if (OFF == 0) { k0 |= f_lo; k1 |= f_hi; }
  else if (OFF < 8) { k0 |= f_lo << (8 * OFF); k1 |= (f_hi << (8 * OFF)) | (f_lo >> (64 - 8 * OFF)); }
  else { k1 |= f_lo << (8 * (OFF - 8)); }
}

// ---- accumulator sinks: where Plan::accumulate sends each (accumulator, value) ---------------------------------------
// Thread-private accumulators in shared memory. Group g owns Plan::PRIV_WORDS * T u64 words:
//   [P8][T] u64   one per live accumulator: the whole value of a one-word accumulator, the LOW word of an i128 sum
//   [P4][T] i32   per i128 sum: the thread's running high-word contribution (value.hi + carry per row: 0 or -1 +- 1
//                 for in-range decimals). A contribution beyond 16 bits goes to the CTA-wide u64 spill word instead, so
//                 with at most 65535 rows per thread (the host sizes the grid for it) the i32 cannot wrap.
// 12 instead of 16 bytes per i128 sum is what lets 640 threads x 5 groups of TPC-H Q1 fit one SM.
template <class Plan>
struct PrivateSink {
  static constexpr bool kSkipAliased = true;
  uint64_t* base;   // &s_acc[g * PRIV_WORDS * T]
  uint64_t* spill;  // &s_spill[g * P4]
  template <int OP, int WORD0, int IDX, int OUT>
  __device__ __forceinline__ void add(const JitArgs&, Val v) {
    constexpr int T = Plan::T;
    uint64_t* p = base + Plan::p8_of(WORD0) * T + threadIdx.x;
    if (OP == OP_ADD_F64) {
      *(double*)p += as_f64(v.lo);
    } else if (OP == OP_ADD_I128) {
      const Val sum = add128(Val{*p, 0}, v);  // sum.hi = v.hi + carry
      *p = sum.lo;
      if (sum.hi != 0) {
        if (sum.hi + 0x8000ULL < 0x10000ULL) {
          uint32_t* ph = (uint32_t*)(base + Plan::P8 * T) + Plan::p4_of(WORD0) * T + threadIdx.x;
          *ph += (uint32_t)sum.hi;
        } else {
          atomicAdd((unsigned long long*)(spill + Plan::p4_of(WORD0)), (unsigned long long)sum.hi);
        }
      }
    } else if (OP == OP_COUNT) {
      p[0] += 1;
    } else {
      p[0] += v.lo;
    }
  }
};

template <int OP>
__device__ __forceinline__ void add_atomic(uint64_t* p, Val v) {
  if (OP == OP_ADD_F64) {
    atomicAdd((double*)p, as_f64(v.lo));
  } else if (OP == OP_ADD_I128) {  // two u64 adds; the carry out of the low word goes to the high word
    const uint64_t old = atomicAdd((unsigned long long*)p, (unsigned long long)v.lo);
    const uint64_t hi = v.hi + ((old + v.lo) < old);
    if (hi) atomicAdd((unsigned long long*)(p + 1), (unsigned long long)hi);
  } else if (OP == OP_COUNT) {
    atomicAdd((unsigned long long*)p, 1ULL);
  } else {
    atomicAdd((unsigned long long*)p, (unsigned long long)v.lo);
  }
}

struct SharedAtomicSink {  // one of `replicas` copies of [n_groups][W] in shared memory
  static constexpr bool kSkipAliased = true;
  uint64_t* base;  // &s_acc[(rep * n_groups + g) * W]
  template <int OP, int WORD0, int IDX, int OUT>
  __device__ __forceinline__ void add(const JitArgs&, Val v) { add_atomic<OP>(base + WORD0, v); }
};

// 64-bit / 128-bit integer adds in SHARED memory as chains of native 32-bit atomics (ATOMS.ADD): the thread whose add
// wraps a limb carries into the next one. (A 64-bit shared atomicAdd compiles to a compare-and-swap spin loop.)
__device__ __forceinline__ void add_shared_limbs(uint32_t* q, const uint32_t* limb, int n) {
  uint32_t carry = 0;
#pragma unroll
  for (int i = 0; i < n; i++) {
    const uint32_t v = limb[i] + carry;
    const uint32_t wrapped = (limb[i] + carry) < carry;  // limb[i] = 0xffffffff and carry = 1: passes the carry on
    carry = wrapped;
    if (v) {
      const uint32_t old = atomicAdd(q + i, v);
      carry += (old + v) < old;
    }
  }
}
struct SharedSplitSink {  // one of `replicas` copies of [g_cap][W] in shared memory
  static constexpr bool kSkipAliased = true;
  uint64_t* base;  // &s_acc[(rep * g_cap + g) * W]
  template <int OP, int WORD0, int IDX, int OUT>
  __device__ __forceinline__ void add(const JitArgs&, Val v) {
    uint32_t* q = (uint32_t*)(base + WORD0);
    if (OP == OP_ADD_F64) {
      atomicAdd((double*)(base + WORD0), as_f64(v.lo));
    } else if (OP == OP_COUNT) {
      const uint32_t old = atomicAdd(q, 1u);
      if (old == 0xffffffffu) atomicAdd(q + 1, 1u);
    } else if (OP == OP_ADD_I128) {
      const uint32_t limb[4] = {(uint32_t)v.lo, (uint32_t)(v.lo >> 32), (uint32_t)v.hi, (uint32_t)(v.hi >> 32)};
      add_shared_limbs(q, limb, 4);
    } else {
      const uint32_t limb[2] = {(uint32_t)v.lo, (uint32_t)(v.lo >> 32)};
      add_shared_limbs(q, limb, 2);
    }
  }
};

// A copy of [g_cap][W] that only ONE lane at a time touches for a given group (warp-private copies, equal group ids
// inside the warp taken in turns): plain read-modify-write. LDS/STS move 128 B per cycle and SM; shared atomics retire
// about one lane every two cycles, and the Float64 one is a compare-and-swap loop.
struct PlainSharedSink {
  static constexpr bool kSkipAliased = true;
  uint64_t* base;  // &s_acc[(warp * g_cap + g) * W]
  template <int OP, int WORD0, int IDX, int OUT>
  __device__ __forceinline__ void add(const JitArgs&, Val v) {
    uint64_t* p = base + WORD0;
    if (OP == OP_ADD_F64) {
      *(double*)p += as_f64(v.lo);
    } else if (OP == OP_ADD_I128) {
      const Val s = add128(Val{p[0], p[1]}, v);
      p[0] = s.lo;
      p[1] = s.hi;
    } else if (OP == OP_COUNT) {
      p[0] += 1;
    } else {
      p[0] += v.lo;
    }
  }
};

// fused MODE 1: fold the CTA's replicas of one accumulator into its slab (one thread per group)
struct FlushReplicas {
  const uint64_t* s_acc;
  uint64_t* slab;
  uint32_t G, replicas;
  int W, T;
  template <int OP, int WORD0, int WORDS, int IDX, bool ALIASED, int OUT>
  __device__ __forceinline__ void acc() {
    if (ALIASED) return;
    for (uint32_t grp = threadIdx.x; grp < G; grp += T) {
      uint64_t* o = slab + (size_t)grp * W + WORD0;
      if (OP == OP_ADD_F64) {
        double f = 0.0;
        for (uint32_t t = 0; t < replicas; t++) f += as_f64(s_acc[((size_t)t * G + grp) * W + WORD0]);
        o[0] = f64_bits(f);
      } else {
        Val s{0, 0};
        for (uint32_t t = 0; t < replicas; t++) {
          const uint64_t* p = &s_acc[((size_t)t * G + grp) * W + WORD0];
          const uint64_t lo = p[0];
          s.lo += lo;
          s.hi += (s.lo < lo);
          if (WORDS > 1) s.hi += p[1];
        }
        o[0] = s.lo;
        if (WORDS > 1) o[1] = s.hi;
      }
    }
  }
};

struct GlobalAtomicSink {
  static constexpr bool kSkipAliased = false;
  uint32_t g;
  template <int OP, int WORD0, int IDX, int OUT>
  __device__ __forceinline__ void add(const JitArgs& a, Val v) {
    add_atomic<OP>(a.t.acc + (size_t)g * a.t.acc_stride + OUT, v);
  }
};

// warp-wide sum of per-lane (lo, hi) partials / doubles
__device__ __forceinline__ Val warp_sum_u128(Val v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const uint64_t lo = __shfl_down_sync(0xffffffffu, v.lo, d);
    const uint64_t hi = __shfl_down_sync(0xffffffffu, v.hi, d);
    v = add128(v, Val{lo, hi});
  }
  return v;
}
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
  return v;
}

__host__ __device__ __forceinline__ size_t fused_slab_words(uint32_t g_cap, int W) {
  return (size_t)g_cap * W + ((size_t)g_cap + 1) / 2;
}

// fused path: reduce the T private copies of one accumulator of every group into the CTA's slab (one warp per group)
template <class Plan>
struct FlushPrivate {
  const uint64_t* s_acc;
  const uint64_t* s_spill;
  uint64_t* slab;
  uint32_t G;
  template <int OP, int WORD0, int WORDS, int IDX, bool ALIASED, int OUT>
  __device__ __forceinline__ void acc() {
    if (ALIASED) return;
    constexpr int T = Plan::T, W = Plan::W;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t grp = warp; grp < G; grp += T / 32) {
      const uint64_t* p = s_acc + ((size_t)grp * Plan::PRIV_WORDS + Plan::p8_of(WORD0)) * T;
      uint64_t* o = slab + (size_t)grp * W + WORD0;
      if (OP == OP_ADD_F64) {
        double f = 0.0;
        for (int th = lane; th < T; th += 32) f += as_f64(p[th]);
        f = warp_sum_f64(f);
        if (lane == 0) o[0] = f64_bits(f);
      } else {
        const int32_t* ph = (const int32_t*)(s_acc + ((size_t)grp * Plan::PRIV_WORDS + Plan::P8) * T) + Plan::p4_of(WORD0) * T;
        Val s{0, 0};
        for (int th = lane; th < T; th += 32) {
          Val x{p[th], 0};
          if (WORDS > 1) x.hi = (uint64_t)(int64_t)ph[th];
          s = add128(s, x);
        }
        s = warp_sum_u128(s);
        if (lane == 0) {
          o[0] = s.lo;
          if (WORDS > 1) o[1] = s.hi + s_spill[(size_t)grp * Plan::P4 + Plan::p4_of(WORD0)];
        }
      }
    }
  }
};

// shared-atomic tier: fold this CTA's replicas into the dense accumulator arrays (one warp per group)
struct FlushShared {
  const uint64_t* s_acc;
  const JitArgs* a;
  uint32_t n_groups, replicas;
  int W, T;
  template <int OP, int WORD0, int WORDS, int IDX, bool ALIASED, int OUT>
  __device__ __forceinline__ void acc() {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t grp = warp; grp < n_groups; grp += T / 32) {
      uint64_t* o = a->t.acc + (size_t)grp * a->t.acc_stride + OUT;
      if (OP == OP_ADD_F64) {
        double f = 0.0;
        for (uint32_t t = lane; t < replicas; t += 32) f += as_f64(s_acc[((size_t)t * n_groups + grp) * W + WORD0]);
        f = warp_sum_f64(f);
        if (lane == 0 && f != 0.0) atomicAdd((double*)o, f);
      } else {
        Val s{0, 0};
        for (uint32_t t = lane; t < replicas; t += 32) {
          const uint64_t* p = &s_acc[((size_t)t * n_groups + grp) * W + WORD0];
          const uint64_t lo = p[0];
          s.lo += lo;
          s.hi += (s.lo < lo);
          if (WORDS > 1) s.hi += p[1];
        }
        s = warp_sum_u128(s);
        // partial counts are sums by now; an aliased count adds its source's partial
        if (lane == 0 && (s.lo | s.hi)) {
          if (OP == OP_ADD_I128) add_atomic<OP_ADD_I128>(o, s);
          else add_atomic<OP_ADD_I64>(o, s);
        }
      }
    }
  }
};

// two-pass TIER 2: fold the CTA's hot-group cache (slot -> group id in s_tag) into the global records, one thread per slot
constexpr uint32_t kEmptyTag = 0xffffffffu;
struct FlushCache {
  const uint64_t* s_acc;
  const uint32_t* s_tag;
  const JitArgs* a;
  uint32_t slots;
  int W, T;
  template <int OP, int WORD0, int WORDS, int IDX, bool ALIASED, int OUT>
  __device__ __forceinline__ void acc() {
    for (uint32_t slot = threadIdx.x; slot < slots; slot += T) {
      const uint32_t grp = s_tag[slot];
      if (grp == kEmptyTag) continue;
      uint64_t* o = a->t.acc + (size_t)grp * a->t.acc_stride + OUT;
      const uint64_t* p = s_acc + (size_t)slot * W + WORD0;
      if (OP == OP_ADD_F64) {
        const double f = as_f64(p[0]);
        if (f != 0.0) atomicAdd((double*)o, f);
      } else {
        const Val v{p[0], WORDS > 1 ? p[1] : 0ULL};
        if (v.lo | v.hi) {
          if (OP == OP_ADD_I128) add_atomic<OP_ADD_I128>(o, v);
          else add_atomic<OP_ADD_I64>(o, v);  // partial counts are sums by now; an aliased count adds its source's partial
        }
      }
    }
  }
};

// ---- kernel bodies ---------------------------------------------------------------------------------------------------
// `Plan` (generated per aggregate and per batch shape by agg_codegen.cu) provides:
//   constants  T (threads), R (rows per thread per tile), W (accumulator words per group), KEY_WORDS, HAS_SEEN
//   struct Raw                                   the raw column values of one row
//   load1<NEED>(a, r, x), load2<NEED>(a, r, x)   independent loads, then the loads that depend on them (utf8 bytes,
//                                                dictionary remap); NEED = bit 0 filter, bit 1 key, bit 2 values
//   filter(a, x) -> bool, key(a, x) -> PackedKey
//   accumulate(a, x, sink, &ovf) -> seen mask    evaluates every aggregate argument and feeds the sink
//   for_each_acc(f)                              f.acc<OP, WORD0, WORDS, IDX, ALIASED>() per accumulator
constexpr int NEED_FILTER = 1, NEED_KEY = 2, NEED_VALUES = 4;

// Shared memory: acc u64[g_cap * W * T] | seen u32[g_cap] | key cache (slots)
// Scratch slab of CTA b: u64[g_cap * W] partial accumulators, then u32[g_cap] seen masks (padded to 8 bytes).
// MODE 0: thread-private accumulators (copies = T; best for a handful of groups).
// MODE 1: `a.replicas` shared copies of [g_cap][W] updated with native 32-bit shared atomics (hundreds to thousands of
//         groups: the accumulators no longer scale with the thread count, so the SM runs at full residency).
template <class Plan, int MODE>
__device__ __forceinline__ void fused_body(const JitArgs& a) {
  constexpr int T = Plan::T, R = Plan::R, W = Plan::W, KW = Plan::KEY_WORDS;
  extern __shared__ __align__(16) uint64_t smem[];
  const uint32_t G = a.g_cap;
  const uint32_t copies = MODE == 0 ? (uint32_t)T : a.replicas;
  uint64_t* s_acc = smem;
  // MODE 0: [G][PRIV_WORDS][T] thread-private words, then the high-word spill u64[G][P4]
  const uint32_t acc_words = MODE == 0 ? G * (uint32_t)(Plan::PRIV_WORDS * T + Plan::P4) : G * W * copies;
  uint64_t* s_spill = s_acc + (size_t)G * Plan::PRIV_WORDS * T;
  uint32_t* s_seen = (uint32_t*)(s_acc + acc_words);
  KeyCache cache{(unsigned char*)(s_seen + ((G + 1) & ~1u)), a.cache_slots - 1};
  constexpr uint32_t CACHE_ENTRY = KW == 1 ? 16 : 24;
  __shared__ int s_abort;
  for (uint32_t i = threadIdx.x; i < acc_words; i += T) s_acc[i] = 0;
  for (uint32_t i = threadIdx.x; i < G; i += T) s_seen[i] = 0;
  for (uint32_t i = threadIdx.x; i < a.cache_slots; i += T) *(uint32_t*)(cache.base + i * CACHE_ENTRY + CACHE_ENTRY - 8) = 0;
  if (threadIdx.x == 0) s_abort = *(volatile int*)&a.t.flags[3];  // an earlier CTA gave up already: the pass is void
  __syncthreads();
  if (s_abort) return;

  bool ovf = false;
  uint32_t tiles_done = 0;
  const uint32_t rep = MODE == 0 ? 0u : (threadIdx.x >> 5) % a.replicas;  // one replica per warp while they last
  const int64_t tile = (int64_t)T * R, last = a.n_rows - 1, stride = (int64_t)gridDim.x * tile;
  // Software pipeline over this CTA's tiles (rows past the end re-read the last row):
  //   tile i+1: bulk L2 prefetch of every streamed buffer; its "early" loads (utf8 offsets, remapped dictionary keys)
  //             are issued into registers while tile i is processed
  //   tile i:   all remaining loads are issued back to back (phase 1), then consumed (phase 2)
  // MODE 1 with one copy per warp: no atomics at all. The lanes of a warp that hold the same group in this step take
  // turns, everything else is a plain read-modify-write of the warp's own copy. Every lane runs the warp-synchronous
  // part, whatever its row did.
  const bool warp_private = MODE == 1 && a.replicas == (uint32_t)(T / 32);
  unsigned char* s_claim = cache.base + (size_t)a.cache_slots * 24;  // u8[T / 32][G rounded up to 16] (host: fused smem)
  auto consume = [&](const typename Plan::Raw& xr, int64_t r) {
    if (MODE == 1 && warp_private) {
      const uint32_t lane = threadIdx.x & 31;
      bool active = r <= last && Plan::filter(a, xr);
      uint32_t gid = 0x80000000u | lane;  // rows that add nothing: a group id of their own
      if (active) {
        const PackedKey k = Plan::key(a, xr);
        gid = a.cache_slots == 0 ? intern<KW>(a.t, k) : cached_intern<KW>(cache, a.t, k);
        if (gid >= G) {
          s_abort = 1;
          active = false;
          gid = 0x80000000u | lane;
        }
      }
      // turns by claiming: every pending lane writes its lane id into the warp's claim byte of its group, the lane that
      // reads its own id back owns the group for this turn. (match.any would give the peers directly, but its cost
      // grows with the number of distinct values in the warp - ~27 of 32 here: measured 2.5 cycles per row and SM.)
      unsigned char* claim = s_claim + (size_t)(threadIdx.x >> 5) * ((G + 15) & ~15u);
      bool pending = active;
      while (__any_sync(0xffffffffu, pending)) {
        if (pending) claim[gid] = (unsigned char)lane;
        __syncwarp();
        if (pending && claim[gid] == (unsigned char)lane) {
          PlainSharedSink sink{s_acc + ((size_t)rep * G + gid) * W};
          const uint32_t seen = Plan::accumulate(a, xr, sink, &ovf);
          if (Plan::HAS_SEEN && seen) atomicOr(&s_seen[gid], seen);
          pending = false;
        }
        __syncwarp();
      }
      return;
    }
    if (r > last || !Plan::filter(a, xr)) return;
    const PackedKey k = Plan::key(a, xr);
    // cache_slots == 0: no per-CTA key cache (all of shared memory holds accumulators; the table is L2 resident)
    const uint32_t gid = (MODE != 0 && a.cache_slots == 0) ? intern<KW>(a.t, k) : cached_intern<KW>(cache, a.t, k);
    if (gid >= G) { s_abort = 1; return; }  // beyond the private capacity (or the global table overflowed)
    uint32_t seen;
    if (MODE == 0) {
      PrivateSink<Plan> sink{s_acc + (size_t)gid * Plan::PRIV_WORDS * T, s_spill + (size_t)gid * Plan::P4};
      seen = Plan::accumulate(a, xr, sink, &ovf);
    } else {
      SharedSplitSink sink{s_acc + ((size_t)rep * G + gid) * W};
      seen = Plan::accumulate(a, xr, sink, &ovf);
    }
    if (Plan::HAS_SEEN && seen) atomicOr(&s_seen[gid], seen);
  };
  const int64_t first = (int64_t)blockIdx.x * tile;
  if (Plan::PIPE) {
    // Three-stage pipeline: while tile i is consumed, every load of tile i+1 is in flight (issued from the early
    // loads of tile i+1, which were requested one iteration ago) and the early loads of tile i+2 are requested.
    // Row indices are Plan::Row (32-bit when the batch allows: one IMAD.WIDE per address).
    typedef typename Plan::Row Row;
    const Row rlast = (Row)last, rstride = (Row)stride;
    typename Plan::Raw0 early[R];
    typename Plan::Raw fa[R], fb[R];  // in flight, being consumed
    Row r0 = (Row)first + threadIdx.x;  // first row of this thread in the tile being consumed
#pragma unroll
    for (int u = 0; u < R; u++) {
      const Row r = r0 + u * T;
      Plan::template load0<7>(a, r < rlast ? r : rlast, early[u]);
    }
    auto issue = [&](typename Plan::Raw (&dst)[R], Row rt) {  // all loads of the tile at rt, early loads of the one after
#pragma unroll
      for (int u = 0; u < R; u++) {
        const Row r = rt + u * T, rc = r < rlast ? r : rlast;
        static_cast<typename Plan::Raw0&>(dst[u]) = early[u];
        Plan::template load1<7>(a, rc, dst[u]);
        Plan::template load2<7>(a, rc, dst[u]);
        const Row rn = r + rstride;
        Plan::template load0<7>(a, rn < rlast ? rn : rlast, early[u]);
      }
    };
    auto stop = [&]() {  // this CTA or (polled every 16th tile) another one gave up: the pass is void
      const bool st = *(volatile int*)&s_abort || ((++tiles_done & 15) == 0 && *(volatile int*)&a.t.flags[3]);
      return MODE == 1 ? (bool)__any_sync(0xffffffffu, st) : st;  // (the warp-private path needs whole warps in the loop)
    };
    issue(fa, r0);
    for (;;) {  // (two tiles per trip with the register sets swapping roles spills at 96 registers: measured 30% slower)
      if (r0 - threadIdx.x > rlast) break;
#pragma unroll
      for (int u = 0; u < R; u++) fb[u] = fa[u];
      issue(fa, r0 + rstride);
#pragma unroll
      for (int u = 0; u < R; u++) consume(fb[u], (int64_t)(r0 + u * T));
      r0 += rstride;
      if (stop()) break;
    }
  } else {
    typename Plan::Raw0 nx[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = first + u * T + threadIdx.x;
      Plan::template load0<7>(a, r < last ? r : last, nx[u]);
    }
    for (int64_t base = first; base < a.n_rows; base += stride) {
      typename Plan::Raw x[R];
      const int64_t next = base + stride;
      if (Plan::L2_PREFETCH && next < a.n_rows)
        Plan::template prefetch_l2<7>(a, next, next + tile <= a.n_rows ? tile : a.n_rows - next);
#pragma unroll
      for (int u = 0; u < R; u++) {
        const int64_t r = base + u * T + threadIdx.x;
        static_cast<typename Plan::Raw0&>(x[u]) = nx[u];
        Plan::template load1<7>(a, r < last ? r : last, x[u]);
        Plan::template load2<7>(a, r < last ? r : last, x[u]);
      }
#pragma unroll
      for (int u = 0; u < R; u++) {
        const int64_t r = next + u * T + threadIdx.x;
        Plan::template load0<7>(a, r < last ? r : last, nx[u]);
      }
#pragma unroll
      for (int u = 0; u < R; u++) consume(x[u], base + u * T + threadIdx.x);
      // some CTA gave up: the pass is void (the global flag is polled every 16th tile: it only saves work)
      const bool st = *(volatile int*)&s_abort || ((++tiles_done & 15) == 0 && *(volatile int*)&a.t.flags[3]);
      if (MODE == 1 ? (bool)__any_sync(0xffffffffu, st) : st) break;
    }
  }
  if (ovf) atomicOr(&a.t.flags[2], 1);
  __syncthreads();
  if (s_abort) {
    if (threadIdx.x == 0) atomicOr(&a.t.flags[3], 1);
    return;
  }
  uint64_t* slab = a.scratch + (size_t)blockIdx.x * fused_slab_words(G, W);
  if (MODE == 0) {
    FlushPrivate<Plan> fl{s_acc, s_spill, slab, G};
    Plan::for_each_acc(fl);
  } else {
    FlushReplicas fl{s_acc, slab, G, a.replicas, W, T};
    Plan::for_each_acc(fl);
  }
  uint32_t* slab_seen = (uint32_t*)(slab + (size_t)G * W);
  for (uint32_t i = threadIdx.x; i < G; i += T) slab_seen[i] = s_seen[i];
}

// two-pass path, pass 0 (only when the batch could overflow the table): linear-counting sketch of the batch's distinct
// keys. Every surviving row sets one bit (chosen by its table hash) of a bitmap of `a.n_groups` 32-bit words
// (a power of two) at `a.gids`; the host turns the fraction of zero bits into a cardinality estimate and sizes the
// table once, instead of discovering the cardinality by overflowing and re-running the intern pass.
template <class Plan>
__device__ __forceinline__ void sketch_body(const JitArgs& a) {
  constexpr int T = Plan::T, R = Plan::R;
  const int64_t tile = (int64_t)T * R, last = a.n_rows - 1;
  const uint32_t wmask = a.n_groups - 1;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < a.n_rows; base += (int64_t)gridDim.x * tile) {
    typename Plan::Raw x[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load0<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
      Plan::template load1<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
    }
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load2<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
    }
    // test before set: with repeated keys the bit is usually there already, and a plain L2 read does not serialise the
    // way an atomic on a hot word does (Zipf(1) keys, 5e8 rows: 25.8 ms -> the uniform-key time)
    uint32_t word[R], bit[R], cur[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      bit[u] = 0;
      word[u] = 0;
      if (r > last || !Plan::filter(a, x[u])) continue;
      const uint64_t h = key_hash<Plan::KEY_WORDS>(Plan::key(a, x[u])) >> 16;  // bits the table index does not use first
      word[u] = (uint32_t)(h >> 5) & wmask;
      bit[u] = 1u << (h & 31);
    }
#pragma unroll
    for (int u = 0; u < R; u++) cur[u] = bit[u] ? __ldcg(&a.gids[word[u]]) : 0u;
#pragma unroll
    for (int u = 0; u < R; u++)
      if (bit[u] & ~cur[u]) atomicOr(&a.gids[word[u]], bit[u]);
  }
}

// two-pass path, radix pre-pass (tables far beyond the L2): the bucket of a row is the top bits of its first slot index, so
// that rows of one bucket touch one contiguous slice of the table; the host scatters the rows into bucket order
// (partition.cu's scan + scatter: the chunk layout and hist[bucket][chunk] are its) before pass 1 and 2 run over them.
// Rows the filter drops go to bucket 0 (they are dropped again later).
template <class Plan>
__device__ __forceinline__ void bucket_body(const JitArgs& a) {
  constexpr int T = Plan::T, R = Plan::R;
  extern __shared__ __align__(16) uint64_t smem[];
  uint32_t* s_hist = (uint32_t*)smem;
  const uint32_t nb = a.n_groups;
  for (uint32_t i = threadIdx.x; i < nb; i += T) s_hist[i] = 0;
  __syncthreads();
  // a.replicas CTAs share one chunk of the scatter's layout (its grid alone would leave the SMs nearly empty here)
  const int64_t tile = (int64_t)T * R, last = a.n_rows - 1;
  const uint32_t chunk = blockIdx.x / a.replicas, sub = blockIdx.x % a.replicas, n_chunks = gridDim.x / a.replicas;
  const int64_t sub_rows = (((int64_t)a.bucket_chunk + a.replicas - 1) / a.replicas + tile - 1) / tile * tile;
  int64_t chunk_end = ((int64_t)chunk + 1) * (int64_t)a.bucket_chunk;
  if (chunk_end > a.n_rows) chunk_end = a.n_rows;
  const int64_t begin = (int64_t)chunk * (int64_t)a.bucket_chunk + (int64_t)sub * sub_rows;
  int64_t end = begin + sub_rows;
  if (end > chunk_end) end = chunk_end;
  for (int64_t base = begin; base < end; base += tile) {
    typename Plan::Raw x[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load0<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
      Plan::template load1<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
    }
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load2<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
    }
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      if (r >= end) continue;
      uint32_t p = 0;
      if (Plan::filter(a, x[u])) p = ((uint32_t)key_hash<Plan::KEY_WORDS>(Plan::key(a, x[u])) & a.t.mask) >> a.bucket_shift;
      a.bucket_ids[r] = (uint16_t)p;
      atomicAdd(&s_hist[p], 1u);
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < nb; i += T)  // (zero-initialised by the host)
    if (s_hist[i]) atomicAdd(&a.bucket_hist[(size_t)i * n_chunks + chunk], s_hist[i]);
}

// two-pass path, pass 1
template <class Plan>
__device__ __forceinline__ void intern_body(const JitArgs& a) {
  constexpr int T = Plan::T, R = Plan::R;
  const int64_t tile = (int64_t)T * R, last = a.n_rows - 1;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < a.n_rows; base += (int64_t)gridDim.x * tile) {
    typename Plan::Raw x[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load0<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
      Plan::template load1<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
    }
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load2<NEED_FILTER | NEED_KEY>(a, r < last ? r : last, x[u]);
    }
    PackedKey k[R];
    bool active[R];
    uint32_t gid[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      active[u] = r <= last && Plan::filter(a, x[u]);
      k[u] = active[u] ? Plan::key(a, x[u]) : PackedKey{0, 0, 0};  // (a dropped row must not raise "key too long")
    }
    warp_intern<Plan::KEY_WORDS, R>(a.t, k, active, gid);
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      if (r <= last) a.gids[r] = gid[u];
    }
  }
}

// two-pass path, pass 2. TIER 1: `replicas` shared-memory copies of [n_groups][W] + atomics; TIER 2: global atomics.
template <class Plan, int TIER>
__device__ __forceinline__ void accumulate_body(const JitArgs& a) {
  constexpr int T = Plan::T, R = Plan::R, W = Plan::W;
  extern __shared__ __align__(16) uint64_t smem[];
  uint64_t* s_acc = smem;
  uint32_t* s_seen = nullptr;
  const uint32_t n_groups = a.n_groups;
  if (TIER == 1) {
    for (uint32_t i = threadIdx.x; i < n_groups * W * a.replicas; i += T) s_acc[i] = 0;
    s_seen = (uint32_t*)(s_acc + (size_t)n_groups * W * a.replicas);
    if (Plan::HAS_SEEN)
      for (uint32_t i = threadIdx.x; i < n_groups; i += T) s_seen[i] = 0;
    __syncthreads();
  }
  // TIER 2 with skewed keys: a handful of groups take a large share of the rows, and their global atomics all land on
  // the same L2 addresses (Zipf(1) over 1e7 groups: the hottest group gets 4 % of the rows; 5e8 rows: 166 ms instead of
  // 43 ms uniform). Each CTA therefore keeps a small direct-mapped cache of groups in shared memory: the first group to
  // arrive at a slot owns it for the launch (a hot group arrives first with high probability), its rows are added with
  // native 32-bit shared atomics, everything else goes to the global records as before; the cache is folded into the
  // global records once at the end.
  uint32_t* s_tag = nullptr;
  uint32_t* s_cseen = nullptr;
  const uint32_t cslots = TIER == 2 ? a.cache_slots : 0;
  if (TIER == 2 && cslots) {
    for (uint32_t i = threadIdx.x; i < cslots * W; i += T) s_acc[i] = 0;
    s_tag = (uint32_t*)(s_acc + (size_t)cslots * W);
    s_cseen = s_tag + cslots;
    for (uint32_t i = threadIdx.x; i < cslots; i += T) {
      s_tag[i] = kEmptyTag;
      if (Plan::HAS_SEEN) s_cseen[i] = 0;
    }
    __syncthreads();
  }
  const uint32_t rep = TIER == 1 ? (threadIdx.x % a.replicas) : 0;
  bool ovf = false;
  const int64_t tile = (int64_t)T * R, last = a.n_rows - 1;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < a.n_rows; base += (int64_t)gridDim.x * tile) {
    typename Plan::Raw x[R];
    uint32_t g[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      g[u] = r <= last ? a.gids[r] : kNoGroup;
      Plan::template load1<NEED_VALUES>(a, r < last ? r : last, x[u]);
    }
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      Plan::template load2<NEED_VALUES>(a, r < last ? r : last, x[u]);
    }
#pragma unroll
    for (int u = 0; u < R; u++) {
      if (g[u] == kNoGroup) continue;
      uint32_t seen;
      if (TIER == 1) {
        SharedAtomicSink sink{s_acc + ((size_t)rep * n_groups + g[u]) * W};
        seen = Plan::accumulate(a, x[u], sink, &ovf);
        if (Plan::HAS_SEEN && seen) atomicOr(&s_seen[g[u]], seen);
      } else {
        bool cached = false;
        uint32_t slot = 0;
        if (cslots) {
          slot = (g[u] * 0x9E3779B1u) >> (32 - __popc(cslots - 1));
          uint32_t tag = s_tag[slot];
          if (tag == kEmptyTag) {
            tag = atomicCAS(&s_tag[slot], kEmptyTag, g[u]);
            if (tag == kEmptyTag) tag = g[u];
          }
          cached = tag == g[u];
        }
        if (cached) {
          SharedSplitSink sink{s_acc + (size_t)slot * W};
          seen = Plan::accumulate(a, x[u], sink, &ovf);
          if (Plan::HAS_SEEN && seen) atomicOr(&s_cseen[slot], seen);
        } else {
          GlobalAtomicSink sink{g[u]};
          seen = Plan::accumulate(a, x[u], sink, &ovf);
          if (Plan::HAS_SEEN && seen) atomicOr(&a.seen[g[u]], seen);
        }
      }
    }
  }
  if (ovf) atomicOr(&a.t.flags[2], 1);
  if (TIER == 2 && cslots) {
    __syncthreads();
    FlushCache fc{s_acc, s_tag, &a, cslots, W, T};
    Plan::for_each_acc(fc);
    if (Plan::HAS_SEEN)
      for (uint32_t i = threadIdx.x; i < cslots; i += T)
        if (s_tag[i] != kEmptyTag && s_cseen[i]) atomicOr(&a.seen[s_tag[i]], s_cseen[i]);
  }
  if (TIER != 1) return;
  __syncthreads();
  FlushShared fl{s_acc, &a, n_groups, a.replicas, W, T};
  Plan::for_each_acc(fl);
  if (Plan::HAS_SEEN)
    for (uint32_t i = threadIdx.x; i < n_groups; i += T)
      if (s_seen[i]) atomicOr(&a.seen[i], s_seen[i]);
}

}  // namespace bgpu

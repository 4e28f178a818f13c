// K4/K5: grouped hash aggregate, Partial and Final modes.
//
// Replaces DataFusion's AggregateExec -> GroupedHashAggregateStream for the aggregates on Ballista's hot path
// (plan shape /root/reference/ballista/scheduler/src/planner.rs:376-427; proto
// /root/reference/ballista/core/proto/datafusion.proto:1405-1411,1437-1450; restated SURVEY.md A.3).
//
// Structure mirrors the reference's stream (intern keys -> dense group ids, then one update per accumulator):
//   phase 1  intern_kernel     : pack each row's group key (<= 16 bytes + null bits), probe/insert an
//                                open-addressing table in global memory, write the dense group id per row.
//                                Idempotent, so a capacity overflow is handled by growing and re-running.
//   phase 2  accumulate_kernel : with the exact group count known, pick a tier:
//              PRIVATE  few groups     : per-thread private accumulators in shared memory, no atomics
//              SHARED   groups fit smem: per-CTA shared accumulators (replicated), shared-memory atomics
//              GLOBAL   otherwise      : atomics straight to the dense accumulator arrays in HBM
//            Int64/Decimal128 sums are wrapping two's-complement adds (u64 adds with carry), so any order is
//            bit-exact; Float64 sums use atomicAdd(double) and carry the north-star tolerance of 1e-9.
//   emit     unpack keys, hand accumulator arrays out as Arrow buffers (state) or evaluate Avg (final).
#include <string.h>

#include <algorithm>

#include "internal.h"

namespace bgpu {

enum ValKind : int32_t { V_NONE = 0, V_I8, V_I16, V_I32, V_I64, V_U8, V_U16, V_U32, V_U64, V_F32, V_F64, V_I128 };
enum AccOp : int32_t { OP_ADD_I64 = 0, OP_ADD_I128 = 1, OP_ADD_F64 = 2, OP_COUNT = 3 };

constexpr int kMaxGroupCols = 8;
constexpr int kMaxAcc = 24;
constexpr int kAggThreads = 256;

struct KeyCol {
  const void* values;
  const int32_t* offsets;
  const uint8_t* validity;
  int32_t kind;      // ColKind
  int32_t width;     // fixed width; utf8: max inline bytes (excluding the length byte)
  int32_t byte_off;  // position inside the 16-byte packed key
};
struct KeyPlan {
  KeyCol cols[kMaxGroupCols];
  int32_t n;
  int32_t words;  // 1: packed key fits 8 bytes, 2: 16 bytes
};

struct GroupTable {
  unsigned long long* entries;  // (tag << 32) | state; state 0 empty, 1 busy, >= 2: gid + 2
  uint32_t mask;                // capacity - 1 (capacity = 2 * max_groups, power of two)
  uint32_t max_groups;
  uint64_t* key0;               // packed keys of group gid (dense)
  uint64_t* key1;
  uint32_t* kmeta;              // null bits of the key columns
  uint32_t* n_groups;           // device counter
  int32_t* flags;               // [0] capacity overflow, [1] string key too long
};

struct AccK {
  const void* values;       // input column (nullptr for COUNT(*))
  const uint8_t* validity;  // nullptr = no nulls in this batch
  uint64_t* out;            // dense accumulator array: words * max_groups u64
  int32_t vk;               // ValKind of the input
  int32_t op;               // AccOp
  int32_t words;            // 1 or 2
  int32_t word0;            // first word of this accumulator inside a group's smem state
  int32_t seen_bit;         // bit in seen[gid]
};
struct AccPlan {
  AccK a[kMaxAcc];
  int32_t n;
  int32_t total_words;
  uint32_t* seen;  // per-group bitmask of accumulators that saw a non-null input; nullptr = not tracked yet
};

// ---- device helpers ---------------------------------------------------------------------------------------
__device__ __forceinline__ bool bit_set(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // murmur3 fmix64: table hash (not the partition hash)
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

struct PackedKey { uint64_t k0, k1; uint32_t meta; };

__device__ __forceinline__ void put_bytes(PackedKey& k, int off, uint64_t v, int nbytes) {
  // place the low nbytes of v at byte offset off of the 128-bit key (fields never straddle the 8-byte boundary
  // except utf8, which is put byte by byte)
  if (off < 8) k.k0 |= v << (8 * off); else k.k1 |= v << (8 * (off - 8));
  (void)nbytes;
}

__device__ __forceinline__ PackedKey pack_key(const KeyPlan& kp, int64_t r, int32_t* flags) {
  PackedKey k{0, 0, 0};
#pragma unroll 1
  for (int c = 0; c < kp.n; c++) {
    const KeyCol& kc = kp.cols[c];
    if (kc.validity && !bit_set(kc.validity, r)) { k.meta |= 1u << c; continue; }
    switch (kc.kind) {
      case K_FIXED:
      case K_DICT32:
        switch (kc.width) {
          case 8: put_bytes(k, kc.byte_off, ((const uint64_t*)kc.values)[r], 8); break;
          case 4: put_bytes(k, kc.byte_off, ((const uint32_t*)kc.values)[r], 4); break;
          case 2: put_bytes(k, kc.byte_off, ((const uint16_t*)kc.values)[r], 2); break;
          case 1: put_bytes(k, kc.byte_off, ((const uint8_t*)kc.values)[r], 1); break;
          default: {  // 16: the only key column
            const ulonglong2 v = ((const ulonglong2*)kc.values)[r];
            k.k0 = v.x; k.k1 = v.y;
          } break;
        }
        break;
      case K_BOOL: put_bytes(k, kc.byte_off, bit_set((const uint8_t*)kc.values, r) ? 1 : 0, 1); break;
      default: {  // K_UTF8 / K_BINARY: length byte then up to `width` bytes
        const int32_t b = kc.offsets[r], len = kc.offsets[r + 1] - b;
        if (len > kc.width) { atomicOr(&flags[1], 1); break; }
        put_bytes(k, kc.byte_off, (uint64_t)len, 1);
        const uint8_t* p = (const uint8_t*)kc.values + b;
        for (int i = 0; i < len; i++) put_bytes(k, kc.byte_off + 1 + i, p[i], 1);
      } break;
    }
  }
  return k;
}

// Find or insert; returns the dense group id (>= max_groups when the table overflowed: the pass is void).
__device__ __forceinline__ uint32_t intern(const GroupTable& t, const PackedKey& k, int words) {
  const uint64_t h = mix64(k.k0 ^ mix64(k.k1 + 0x9e3779b97f4a7c15ULL * (k.meta + 1)));
  const unsigned long long tag = (h >> 32) << 32;
  uint32_t i = (uint32_t)h & t.mask;
  while (true) {
    unsigned long long e = *(volatile unsigned long long*)&t.entries[i];
    if (e == 0) {
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
}

__global__ void __launch_bounds__(kAggThreads)
    intern_kernel(const __grid_constant__ KeyPlan kp, const __grid_constant__ GroupTable t, int64_t n_rows,
                  uint32_t* __restrict__ gid_out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    const PackedKey k = pack_key(kp, r, t.flags);
    gid_out[r] = intern(t, k, kp.words);
  }
}

// Rebuild the entry table from the dense keys (after growing).
__global__ void __launch_bounds__(kAggThreads) rehash_kernel(const __grid_constant__ GroupTable t, uint32_t n, int words) {
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) {
    const uint64_t k0 = t.key0[g], k1 = words > 1 ? t.key1[g] : 0;
    const uint32_t meta = t.kmeta[g];
    const uint64_t h = mix64(k0 ^ mix64(k1 + 0x9e3779b97f4a7c15ULL * (meta + 1)));
    const unsigned long long tag = (h >> 32) << 32;
    uint32_t i = (uint32_t)h & t.mask;
    while (atomicCAS(&t.entries[i], 0ULL, tag | (unsigned long long)(g + 2u)) != 0ULL) i = (i + 1) & t.mask;
  }
}

// ---- phase 2: accumulate -------------------------------------------------------------------------------------
struct Val { uint64_t lo, hi; };  // i128 (sign-extended) or the bits of a double in lo

__device__ __forceinline__ Val load_val(const AccK& a, int64_t r) {
  Val v{1, 0};
  switch (a.vk) {
    case V_I64: case V_U64: v.lo = ((const uint64_t*)a.values)[r]; break;
    case V_I128: { const ulonglong2 x = ((const ulonglong2*)a.values)[r]; v.lo = x.x; v.hi = x.y; } break;
    case V_F64: v.lo = ((const uint64_t*)a.values)[r]; break;
    case V_I32: v.lo = (uint64_t)(int64_t)((const int32_t*)a.values)[r]; break;
    case V_U32: v.lo = ((const uint32_t*)a.values)[r]; break;
    case V_F32: v.lo = (uint64_t)__double_as_longlong((double)((const float*)a.values)[r]); break;
    case V_I16: v.lo = (uint64_t)(int64_t)((const int16_t*)a.values)[r]; break;
    case V_U16: v.lo = ((const uint16_t*)a.values)[r]; break;
    case V_I8: v.lo = (uint64_t)(int64_t)((const int8_t*)a.values)[r]; break;
    case V_U8: v.lo = ((const uint8_t*)a.values)[r]; break;
    default: break;  // V_NONE: COUNT(*) adds 1
  }
  if (a.op == OP_ADD_F64 && a.vk != V_F64 && a.vk != V_F32) {  // planner coercion: avg(int) -> avg(CAST AS Float64)
    const bool is_unsigned = a.vk == V_U8 || a.vk == V_U16 || a.vk == V_U32 || a.vk == V_U64;
    v.lo = (uint64_t)__double_as_longlong(is_unsigned ? (double)v.lo : (double)(int64_t)v.lo);
  }
  if (a.op == OP_COUNT) v.lo = 1;
  return v;
}

enum Tier : int { TIER_PRIVATE = 0, TIER_SHARED = 1, TIER_GLOBAL = 2 };

// plain (non-atomic) add into a thread-private slot; `stride` = distance between the words of one accumulator
__device__ __forceinline__ void add_private(uint64_t* p, int stride, const AccK& a, const Val& v) {
  if (a.op == OP_ADD_F64) {
    *(double*)p += __longlong_as_double((long long)v.lo);
  } else if (a.op == OP_ADD_I128) {
    const uint64_t old = p[0], lo = old + v.lo;
    p[0] = lo;
    p[stride] += v.hi + (lo < old);
  } else {
    p[0] += v.lo;
  }
}
// atomic add (shared or global memory); i128 = two u64 adds, the carry out of the low word goes to the high word
__device__ __forceinline__ void add_atomic(uint64_t* p, const AccK& a, const Val& v) {
  if (a.op == OP_ADD_F64) {
    atomicAdd((double*)p, __longlong_as_double((long long)v.lo));
  } else if (a.op == OP_ADD_I128) {
    const uint64_t old = atomicAdd((unsigned long long*)p, (unsigned long long)v.lo);
    const uint64_t hi = v.hi + ((old + v.lo) < old);
    if (hi) atomicAdd((unsigned long long*)(p + 1), (unsigned long long)hi);
  } else {
    atomicAdd((unsigned long long*)p, (unsigned long long)v.lo);
  }
}

constexpr int kRowsPerIter = 4;

template <int TIER>
__global__ void __launch_bounds__(kAggThreads)
    accumulate_kernel(const __grid_constant__ AccPlan plan, const uint32_t* __restrict__ gids, int64_t n_rows,
                      uint32_t n_groups, int replicas) {
  extern __shared__ __align__(16) uint64_t s_acc[];
  const int W = plan.total_words;
  const int T = kAggThreads;
  uint32_t* s_seen = nullptr;
  if (TIER == TIER_PRIVATE) {
    for (uint32_t i = threadIdx.x; i < n_groups * W * T; i += T) s_acc[i] = 0;
    s_seen = (uint32_t*)(s_acc + (size_t)n_groups * W * T);
  } else if (TIER == TIER_SHARED) {
    for (uint32_t i = threadIdx.x; i < n_groups * W * replicas; i += T) s_acc[i] = 0;
    s_seen = (uint32_t*)(s_acc + (size_t)n_groups * W * replicas);
  }
  if (TIER != TIER_GLOBAL && plan.seen)
    for (uint32_t i = threadIdx.x; i < n_groups; i += T) s_seen[i] = 0;
  if (TIER != TIER_GLOBAL) __syncthreads();
  const int rep = TIER == TIER_SHARED ? (threadIdx.x % replicas) : 0;

  const int64_t step = (int64_t)gridDim.x * T * kRowsPerIter;
  for (int64_t base = (int64_t)blockIdx.x * T * kRowsPerIter; base < n_rows; base += step) {
    uint32_t g[kRowsPerIter];
    uint32_t seen_mask[kRowsPerIter];
#pragma unroll
    for (int u = 0; u < kRowsPerIter; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      g[u] = r < n_rows ? gids[r] : 0xffffffffu;
      seen_mask[u] = 0;
    }
#pragma unroll 1
    for (int ai = 0; ai < plan.n; ai++) {
      const AccK& a = plan.a[ai];
      Val v[kRowsPerIter];
      bool ok[kRowsPerIter];
#pragma unroll
      for (int u = 0; u < kRowsPerIter; u++) {
        const int64_t r = base + u * T + threadIdx.x;
        ok[u] = g[u] != 0xffffffffu && (!a.validity || bit_set(a.validity, r));
        if (ok[u]) v[u] = load_val(a, r);
      }
#pragma unroll
      for (int u = 0; u < kRowsPerIter; u++) {
        if (!ok[u]) continue;
        seen_mask[u] |= 1u << a.seen_bit;
        if (TIER == TIER_PRIVATE) {
          add_private(&s_acc[((size_t)g[u] * W + a.word0) * T + threadIdx.x], T, a, v[u]);
        } else if (TIER == TIER_SHARED) {
          add_atomic(&s_acc[((size_t)rep * n_groups + g[u]) * W + a.word0], a, v[u]);
        } else {
          add_atomic(&a.out[(size_t)g[u] * a.words], a, v[u]);
        }
      }
    }
    if (plan.seen) {
#pragma unroll
      for (int u = 0; u < kRowsPerIter; u++) {
        if (g[u] == 0xffffffffu || !seen_mask[u]) continue;
        if (TIER == TIER_GLOBAL) atomicOr(&plan.seen[g[u]], seen_mask[u]);
        else atomicOr(&s_seen[g[u]], seen_mask[u]);
      }
    }
  }
  if (TIER == TIER_GLOBAL) return;
  __syncthreads();
  // flush this CTA's partial state: one warp per (group, accumulator)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = T / 32;
  for (uint32_t item = warp; item < n_groups * (uint32_t)plan.n; item += n_warps) {
    const uint32_t grp = item / plan.n;
    const AccK& a = plan.a[item % plan.n];
    unsigned __int128 isum = 0;
    double fsum = 0.0;
    const int n_parts = TIER == TIER_PRIVATE ? T : replicas;
    for (int t = lane; t < n_parts; t += 32) {
      const uint64_t* p = TIER == TIER_PRIVATE ? &s_acc[((size_t)grp * W + a.word0) * T + t]
                                               : &s_acc[((size_t)t * n_groups + grp) * W + a.word0];
      const int stride = TIER == TIER_PRIVATE ? T : 1;
      if (a.op == OP_ADD_F64) fsum += *(const double*)p;
      else if (a.op == OP_ADD_I128) isum += ((unsigned __int128)p[stride] << 64) | p[0];
      else isum += p[0];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      if (a.op == OP_ADD_F64) {
        fsum += __shfl_down_sync(0xffffffffu, fsum, d);
      } else {
        const uint64_t lo = __shfl_down_sync(0xffffffffu, (uint64_t)isum, d);
        const uint64_t hi = __shfl_down_sync(0xffffffffu, (uint64_t)(isum >> 64), d);
        isum += ((unsigned __int128)hi << 64) | lo;
      }
    }
    if (lane == 0) {
      Val v{(uint64_t)isum, (uint64_t)(isum >> 64)};
      if (a.op == OP_ADD_F64) v.lo = (uint64_t)__double_as_longlong(fsum);
      const bool nonzero = a.op == OP_ADD_F64 ? fsum != 0.0 : (isum != 0);
      if (nonzero) add_atomic(&a.out[(size_t)grp * a.words], a, v);
    }
  }
  if (plan.seen)
    for (uint32_t i = threadIdx.x; i < n_groups; i += T)
      if (s_seen[i]) atomicOr(&plan.seen[i], s_seen[i]);
}

// ---- emit kernels -----------------------------------------------------------------------------------------------
// Unpack one fixed-width / bool / dictionary key column of the dense key arrays into an Arrow column.
__global__ void __launch_bounds__(kAggThreads)
    unpack_fixed_kernel(const uint64_t* __restrict__ key0, const uint64_t* __restrict__ key1,
                        const uint32_t* __restrict__ kmeta, uint32_t n, int col, int kind, int width, int byte_off,
                        void* __restrict__ out_values, uint32_t* __restrict__ out_valid) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false, bit = false;
    if (g < n) {
      valid = !((kmeta[g] >> col) & 1);
      const uint64_t k0 = key0[g], k1 = key1 ? key1[g] : 0;
      const uint64_t word = byte_off < 8 ? (k0 >> (8 * byte_off)) : (k1 >> (8 * (byte_off - 8)));
      bit = (word & 1) != 0;
      if (kind == K_BOOL) {
        // bit-packed output: written by the ballot below
      } else if (width == 16) {
        ((ulonglong2*)out_values)[g] = make_ulonglong2(k0, k1);
      } else if (width == 8) ((uint64_t*)out_values)[g] = word;
      else if (width == 4) ((uint32_t*)out_values)[g] = (uint32_t)word;
      else if (width == 2) ((uint16_t*)out_values)[g] = (uint16_t)word;
      else ((uint8_t*)out_values)[g] = (uint8_t)word;
    }
    if (kind == K_BOOL) {
      const uint32_t bits = __ballot_sync(0xffffffffu, bit);
      if ((threadIdx.x & 31) == 0) ((uint32_t*)out_values)[g >> 5] = bits;
    }
    if (out_valid) {
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
    }
  }
}

// utf8 key column: lengths (pass 0) or bytes (pass 1, offsets known)
__global__ void __launch_bounds__(kAggThreads)
    unpack_utf8_kernel(const uint64_t* __restrict__ key0, const uint64_t* __restrict__ key1,
                       const uint32_t* __restrict__ kmeta, uint32_t n, int col, int byte_off, int pass,
                       int32_t* __restrict__ lens_or_offsets, uint8_t* __restrict__ out_data,
                       uint32_t* __restrict__ out_valid) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false;
    if (g < n) {
      valid = !((kmeta[g] >> col) & 1);
      const uint64_t k0 = key0[g], k1 = key1 ? key1[g] : 0;
      auto byte_at = [&](int off) -> uint32_t {
        return (uint32_t)((off < 8 ? (k0 >> (8 * off)) : (k1 >> (8 * (off - 8)))) & 0xff);
      };
      const int len = valid ? (int)byte_at(byte_off) : 0;
      if (pass == 0) {
        lens_or_offsets[g] = len;
      } else {
        const int32_t b = lens_or_offsets[g];
        for (int i = 0; i < len; i++) out_data[b + i] = (uint8_t)byte_at(byte_off + 1 + i);
      }
    }
    if (pass == 0 && out_valid) {
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
    }
  }
}

// validity bitmap of an accumulator from the seen masks
__global__ void __launch_bounds__(kAggThreads)
    seen_bitmap_kernel(const uint32_t* __restrict__ seen, uint32_t n, int bit, uint32_t* __restrict__ out_valid) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    const uint32_t vb = __ballot_sync(0xffffffffu, g < n && ((seen[g] >> bit) & 1));
    if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
  }
}

__global__ void __launch_bounds__(kAggThreads) fill_u32_kernel(uint32_t* p, uint32_t n, uint32_t v) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = v;
}

// AvgGroupsAccumulator::evaluate. Float64: sum / count. Decimal128: (sum * 10^scale_up) / count, truncating;
// an overflowing scaled multiply is an error in the reference -> flags[2].
__global__ void __launch_bounds__(kAggThreads)
    avg_evaluate_kernel(const uint64_t* __restrict__ sum, const uint64_t* __restrict__ count, uint32_t n, int is_decimal,
                        int scale_up, uint64_t* __restrict__ out, uint32_t* __restrict__ out_valid, int32_t* flags) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false;
    if (g < n) {
      const uint64_t c = count[g];
      valid = c != 0;
      if (is_decimal) {
        __int128 s = (__int128)(((unsigned __int128)sum[2 * g + 1] << 64) | sum[2 * g]);
        __int128 q = 0;
        if (valid) {
          bool ovf = false;
          for (int i = 0; i < scale_up; i++) {
            const __int128 lim = (((__int128)1 << 126) - 1 + ((__int128)1 << 126)) / 10;  // i128::MAX / 10
            if (s > lim || s < -lim) ovf = true;
            s *= 10;
          }
          if (ovf) atomicOr(&flags[2], 1);
          q = s / (__int128)c;
        }
        out[2 * g] = (uint64_t)q;
        out[2 * g + 1] = (uint64_t)((unsigned __int128)q >> 64);
      } else {
        const double s = __longlong_as_double((long long)sum[g]);
        out[g] = (uint64_t)__double_as_longlong(valid ? s / (double)c : 0.0);
      }
    }
    const uint32_t vb = __ballot_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
  }
}
}  // namespace bgpu

using namespace bgpu;

namespace {
struct PhysAcc {
  int32_t op = 0;
  int32_t in_col = -1;  // column of the batch passed to update(); -1 = COUNT(*)
  int32_t vk = V_NONE;
  int32_t words = 1;
  int32_t word0 = 0;
  DevAllocPtr buf;  // words * max_groups u64
};
struct LogicalAgg {
  int32_t fn = 0;
  int32_t phys[2] = {-1, -1};  // Sum: [sum]; Count: [count]; Avg: [count, sum]
  std::string name;
  std::string sum_format, result_format;  // Arrow formats of the state sum and of the final value
  bool decimal = false;
  int scale_up = 0;  // Avg(decimal): result scale - sum scale
};

ValKind vk_of(const Column& c) {
  if (c.kind != K_FIXED) return V_NONE;
  switch (c.format[0]) {
    case 'c': return V_I8;  case 'C': return V_U8;  case 's': return V_I16; case 'S': return V_U16;
    case 'i': return V_I32; case 'I': return V_U32; case 'l': return V_I64; case 'L': return V_U64;
    case 'f': return V_F32; case 'g': return V_F64; case 'd': return V_I128;
    default: return V_NONE;
  }
}
bool is_float(int vk) { return vk == V_F32 || vk == V_F64; }
bool is_unsigned(int vk) { return vk == V_U8 || vk == V_U16 || vk == V_U32 || vk == V_U64; }

uint32_t next_pow2(uint64_t x) {
  uint64_t p = 1;
  while (p < x) p <<= 1;
  return (uint32_t)std::min<uint64_t>(p, 1ULL << 31);
}
}  // namespace

struct bgpu_agg {
  bgpu_ctx* ctx = nullptr;
  int32_t mode = 0;
  bool final_mode = false;
  std::vector<Column> fields;       // raw input schema
  std::vector<int32_t> group_cols;  // indices into `fields`
  std::vector<int32_t> key_off, key_width;  // packed-key layout per group column
  int32_t key_words = 1;
  std::vector<PhysAcc> phys;
  std::vector<LogicalAgg> aggs;
  int32_t total_words = 0;
  // group table
  uint32_t max_groups = 0, n_groups = 0;
  DevAllocPtr entries, key0, key1, kmeta, counters /* [0]=n_groups, [1..3]=flags */, seen;
  bool seen_tracking = false;
  std::vector<std::shared_ptr<Column>> dicts;  // dictionary of each dictionary-typed group column
  std::vector<std::vector<DevAllocPtr>> dict_owned;
  bool emitted = false;
  uint32_t* d_n_groups() { return (uint32_t*)counters->p; }
  int32_t* d_flags() { return (int32_t*)counters->p + 1; }
};

static int agg_launch_grid(bgpu_ctx* ctx, int64_t n, int per_block) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, per_block), (int64_t)ctx->sm_count * 8));
}

// (Re)allocate the table and accumulators for `new_max` groups, keeping the first `keep` groups.
static int agg_resize(bgpu_agg* g, uint32_t new_max, uint32_t keep) {
  bgpu_ctx* ctx = g->ctx;
  DevAllocPtr entries, key0, key1, kmeta, seen;
  BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 2 * 8, &entries));
  BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 8, &key0));
  if (g->key_words > 1) BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 8, &key1));
  BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 4, &kmeta));
  BGPU_CUDA(ctx, cudaMemsetAsync(entries->p, 0, (size_t)new_max * 2 * 8, ctx->stream));
  if (keep) {
    BGPU_CUDA(ctx, cudaMemcpyAsync(key0->p, g->key0->p, (size_t)keep * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    if (key1) BGPU_CUDA(ctx, cudaMemcpyAsync(key1->p, g->key1->p, (size_t)keep * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(kmeta->p, g->kmeta->p, (size_t)keep * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  }
  for (PhysAcc& p : g->phys) {
    DevAllocPtr nb;
    const size_t bytes = (size_t)new_max * p.words * 8;
    BGPU_TRY(dev_alloc(ctx, bytes, &nb));
    const size_t keep_bytes = (size_t)keep * p.words * 8;
    if (keep) BGPU_CUDA(ctx, cudaMemcpyAsync(nb->p, p.buf->p, keep_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemsetAsync((char*)nb->p + keep_bytes, 0, bytes - keep_bytes, ctx->stream));
    p.buf = nb;
  }
  if (g->seen_tracking) {
    BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 4, &seen));
    if (keep) BGPU_CUDA(ctx, cudaMemcpyAsync(seen->p, g->seen->p, (size_t)keep * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemsetAsync((char*)seen->p + (size_t)keep * 4, 0, (size_t)(new_max - keep) * 4, ctx->stream));
    g->seen = seen;
  }
  g->entries = entries; g->key0 = key0; g->key1 = key1; g->kmeta = kmeta;
  g->max_groups = new_max;
  if (keep) {
    GroupTable t{(unsigned long long*)entries->p, new_max * 2 - 1, new_max, (uint64_t*)key0->p,
                 key1 ? (uint64_t*)key1->p : nullptr, (uint32_t*)kmeta->p, g->d_n_groups(), g->d_flags()};
    rehash_kernel<<<agg_launch_grid(ctx, keep, kAggThreads), kAggThreads, 0, ctx->stream>>>(t, keep, g->key_words);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
  }
  return BGPU_OK;
}

static std::string decimal_format(int p, int s) { return "d:" + std::to_string(p) + "," + std::to_string(s); }

extern "C" int bgpu_agg_create(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                               const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                               const int32_t* agg_cols, const char* const* agg_names, int32_t n_agg, bgpu_agg** out) {
  if (!ctx || !out) return BGPU_EINVAL;
  *out = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (!input_schema || !input_schema->format || strcmp(input_schema->format, "+s") != 0)
    return ctx->fail(BGPU_EINVAL, "input_schema must be a struct schema");
  if (mode < BGPU_AGG_MODE_PARTIAL || mode > BGPU_AGG_MODE_SINGLE) return ctx->fail(BGPU_EINVAL, "bad aggregate mode");
  if (n_group < 1 || n_group > kMaxGroupCols)
    return ctx->fail(BGPU_EUNSUPPORTED, "1..8 group columns supported (no-GROUP-BY aggregates use AggregateStream)");
  if (n_agg < 0 || n_agg > 16) return ctx->fail(BGPU_EUNSUPPORTED, "at most 16 aggregates");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  std::unique_ptr<bgpu_agg> g(new bgpu_agg());
  g->ctx = ctx;
  g->mode = mode;
  g->final_mode = mode == BGPU_AGG_MODE_FINAL || mode == BGPU_AGG_MODE_FINAL_PARTITIONED;
  g->fields.resize((size_t)input_schema->n_children);
  for (int64_t i = 0; i < input_schema->n_children; i++)
    BGPU_TRY(parse_field(ctx, input_schema->children[i], &g->fields[(size_t)i]));
  // packed key layout: fields in order, never straddling the 8-byte boundary; utf8 shares what is left
  int fixed_bytes = 0, n_var = 0;
  for (int32_t i = 0; i < n_group; i++) {
    const int32_t c = group_cols[i];
    if (c < 0 || c >= (int32_t)g->fields.size()) return ctx->fail(BGPU_EINVAL, "group column out of range");
    const Column& f = g->fields[(size_t)c];
    g->group_cols.push_back(c);
    if (f.kind == K_UTF8 || f.kind == K_BINARY) n_var++;
    else fixed_bytes += f.kind == K_BOOL ? 1 : f.width;
  }
  int off = 0;
  const int var_budget = n_var ? (16 - fixed_bytes) / n_var : 0;  // bytes per utf8 key incl. its length byte
  if (fixed_bytes > 16 || (n_var && var_budget < 2))
    return ctx->fail(BGPU_EUNSUPPORTED, "group key wider than 16 packed bytes");
  for (int32_t i = 0; i < n_group; i++) {
    const Column& f = g->fields[(size_t)g->group_cols[(size_t)i]];
    const bool var = f.kind == K_UTF8 || f.kind == K_BINARY;
    const int w = var ? var_budget : (f.kind == K_BOOL ? 1 : f.width);
    if (!var && off < 8 && off + w > 8) off = 8;  // do not straddle
    if (off + w > 16) return ctx->fail(BGPU_EUNSUPPORTED, "group key wider than 16 packed bytes");
    g->key_off.push_back(off);
    g->key_width.push_back(var ? w - 1 : w);
    off += w;
  }
  g->key_words = off > 8 ? 2 : 1;
  g->dicts.resize((size_t)n_group);
  g->dict_owned.resize((size_t)n_group);

  // accumulators (deduplicated: the reference computes sum(x) and avg(x)'s sum separately; the values are equal)
  auto add_phys = [&](int32_t op, int32_t in_col, int32_t vk) {
    for (size_t i = 0; i < g->phys.size(); i++)
      if (g->phys[i].op == op && g->phys[i].in_col == in_col && g->phys[i].vk == vk) return (int32_t)i;
    PhysAcc p;
    p.op = op; p.in_col = in_col; p.vk = vk; p.words = op == OP_ADD_I128 ? 2 : 1;
    g->phys.push_back(p);
    return (int32_t)g->phys.size() - 1;
  };
  int32_t state_col = n_group;  // Final modes: next state column of the incoming batches
  for (int32_t i = 0; i < n_agg; i++) {
    LogicalAgg a;
    a.fn = agg_fns[i];
    a.name = (agg_names && agg_names[i]) ? agg_names[i] : ("agg" + std::to_string(i));
    const int32_t c = agg_cols[i];
    if (c >= (int32_t)g->fields.size()) return ctx->fail(BGPU_EINVAL, "aggregate column out of range");
    const Column* f = c >= 0 ? &g->fields[(size_t)c] : nullptr;
    const int32_t vk = f ? vk_of(*f) : V_NONE;
    int p = 0, s = 0;
    if (vk == V_I128) sscanf(f->format.c_str(), "d:%d,%d", &p, &s);
    if (a.fn == BGPU_AGG_COUNT) {
      a.result_format = "l";
      a.phys[0] = g->final_mode ? add_phys(OP_ADD_I64, state_col++, V_I64) : add_phys(OP_COUNT, c, V_NONE);
    } else if (a.fn == BGPU_AGG_SUM || a.fn == BGPU_AGG_AVG) {
      if (!f || vk == V_NONE) return ctx->fail(BGPU_EUNSUPPORTED, "SUM/AVG need a numeric or Decimal128 argument");
      const bool avg = a.fn == BGPU_AGG_AVG;
      int32_t op, state_vk;
      if (vk == V_I128) {
        op = OP_ADD_I128; state_vk = V_I128; a.decimal = true;
        a.sum_format = decimal_format(std::min(38, p + 10), s);
        a.result_format = avg ? decimal_format(std::min(38, p + 4), std::min(38, s + 4)) : a.sum_format;
        a.scale_up = std::min(38, s + 4) - s;
      } else if (avg || is_float(vk)) {
        op = OP_ADD_F64; state_vk = V_F64; a.sum_format = a.result_format = "g";
      } else {
        op = OP_ADD_I64; state_vk = is_unsigned(vk) ? V_U64 : V_I64;
        a.sum_format = a.result_format = is_unsigned(vk) ? "L" : "l";
      }
      if (avg) a.phys[0] = g->final_mode ? add_phys(OP_ADD_I64, state_col++, V_U64) : add_phys(OP_COUNT, c, V_NONE);
      const int32_t sum_phys = g->final_mode ? add_phys(op, state_col++, state_vk) : add_phys(op, c, vk);
      a.phys[avg ? 1 : 0] = sum_phys;
    } else {
      return ctx->fail(BGPU_EUNSUPPORTED, "aggregate function not on the hot path (SUM, AVG, COUNT are)");
    }
    g->aggs.push_back(a);
  }
  if ((int)g->phys.size() > kMaxAcc) return ctx->fail(BGPU_EUNSUPPORTED, "too many accumulators");
  for (PhysAcc& p : g->phys) { p.word0 = g->total_words; g->total_words += p.words; }
  BGPU_TRY(dev_alloc(ctx, 16, &g->counters));
  BGPU_CUDA(ctx, cudaMemsetAsync(g->counters->p, 0, 16, ctx->stream));
  BGPU_TRY(agg_resize(g.get(), 1u << 16, 0));
  *out = g.release();
  return BGPU_OK;
}

extern "C" void bgpu_agg_destroy(bgpu_agg* g) {
  if (!g) return;
  if (g->ctx) cudaSetDevice(g->ctx->device);
  delete g;
}

static bool same_shape(const Column& a, const Column& b) { return a.kind == b.kind && a.width == b.width; }

extern "C" int bgpu_agg_update(bgpu_agg* g, const bgpu_batch* batch) {
  if (!g || !batch) return BGPU_EINVAL;
  bgpu_ctx* ctx = g->ctx;
  if (g->emitted) return ctx->fail(BGPU_EINVAL, "aggregate already emitted");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t n_rows = batch->n_rows;
  const int32_t n_group = (int32_t)g->group_cols.size();
  if (n_rows > 0xfffffff0LL) return ctx->fail(BGPU_EUNSUPPORTED, "more than 2^32 rows in one batch");
  // locate key columns in this batch
  KeyPlan kp{};
  kp.n = n_group;
  kp.words = g->key_words;
  for (int32_t i = 0; i < n_group; i++) {
    const int32_t bc = g->final_mode ? i : g->group_cols[(size_t)i];
    if (bc >= (int32_t)batch->cols.size()) return ctx->fail(BGPU_EINVAL, "batch has too few columns");
    const Column& c = batch->cols[(size_t)bc];
    const Column& f = g->fields[(size_t)g->group_cols[(size_t)i]];
    if (!same_shape(c, f)) return ctx->fail(BGPU_EINVAL, "group column " + f.name + " does not match input_schema");
    if (c.kind == K_DICT32) {
      if (!g->dicts[(size_t)i]) {
        g->dicts[(size_t)i] = c.dict;
        g->dict_owned[(size_t)i] = c.owned;
      } else if (g->dicts[(size_t)i]->values != c.dict->values || g->dicts[(size_t)i]->length != c.dict->length) {
        return ctx->fail(BGPU_EUNSUPPORTED, "dictionary of group column " + f.name + " changed between batches");
      }
    }
    kp.cols[i] = KeyCol{c.values, c.offsets, c.validity, c.kind, g->key_width[(size_t)i], g->key_off[(size_t)i]};
  }
  // accumulator inputs
  AccPlan plan{};
  plan.n = (int32_t)g->phys.size();
  plan.total_words = g->total_words;
  bool any_validity = false;
  for (int32_t i = 0; i < plan.n; i++) {
    const PhysAcc& p = g->phys[(size_t)i];
    AccK& a = plan.a[i];
    a = AccK{nullptr, nullptr, nullptr, p.vk, p.op, p.words, p.word0, i};
    if (p.in_col >= 0) {
      if (p.in_col >= (int32_t)batch->cols.size()) return ctx->fail(BGPU_EINVAL, "batch has too few columns");
      const Column& c = batch->cols[(size_t)p.in_col];
      if (p.op != OP_COUNT) {
        const int need = p.vk == V_I128 ? 16 : (p.vk == V_I64 || p.vk == V_U64 || p.vk == V_F64) ? 8
                         : (p.vk == V_I32 || p.vk == V_U32 || p.vk == V_F32) ? 4 : (p.vk == V_I16 || p.vk == V_U16) ? 2 : 1;
        if (c.kind != K_FIXED || c.width != need)
          return ctx->fail(BGPU_EINVAL, "aggregate input column " + c.name + " has an unexpected type");
        a.values = c.values;
      }
      a.validity = c.validity;
      any_validity |= c.validity != nullptr;
    }
  }
  if (n_rows == 0) return BGPU_OK;

  DevAllocPtr gids;
  BGPU_TRY(dev_alloc(ctx, (size_t)n_rows * 4, &gids));
  const uint32_t prev_groups = g->n_groups;
  uint32_t* h = (uint32_t*)ctx->pinned;
  for (int attempt = 0;; attempt++) {
    GroupTable t{(unsigned long long*)g->entries->p, g->max_groups * 2 - 1, g->max_groups, (uint64_t*)g->key0->p,
                 g->key1 ? (uint64_t*)g->key1->p : nullptr, (uint32_t*)g->kmeta->p, g->d_n_groups(), g->d_flags()};
    const int grid = (int)std::min<int64_t>(ceil_div64(n_rows, kAggThreads), (int64_t)ctx->sm_count * 8);
    intern_kernel<<<grid, kAggThreads, 0, ctx->stream>>>(kp, t, n_rows, (uint32_t*)gids->p);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
    BGPU_CUDA(ctx, cudaMemcpyAsync(h, g->counters->p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint32_t counter = h[0];
    if (h[2]) return ctx->fail(BGPU_EUNSUPPORTED, "string group key longer than the packed-key budget (" +
                                                      std::to_string(16 / std::max(1, n_group) - 1) + " bytes)");
    if (!h[1]) {
      g->n_groups = counter;
      break;
    }
    if (attempt >= 4 || counter > 0x70000000u) return ctx->fail(BGPU_EUNSUPPORTED, "too many groups");
    // overflow: groups [0, max_groups) are intact; size for the distinct count just observed and run again
    const uint32_t keep = g->max_groups;
    const uint32_t new_max = next_pow2(std::max<uint64_t>((uint64_t)counter + counter / 4, (uint64_t)keep * 2));
    const uint32_t reset[4] = {keep, 0, 0, 0};
    BGPU_CUDA(ctx, cudaMemcpyAsync(g->counters->p, reset, 16, cudaMemcpyHostToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    BGPU_TRY(agg_resize(g, new_max, keep));
  }

  if (any_validity && !g->seen_tracking) {
    // first batch with nulls in an aggregate input: groups created by earlier (null-free) batches have seen
    // every accumulator; groups new in this batch start at 0
    BGPU_TRY(dev_alloc(ctx, (size_t)g->max_groups * 4, &g->seen));
    BGPU_CUDA(ctx, cudaMemsetAsync(g->seen->p, 0, (size_t)g->max_groups * 4, ctx->stream));
    if (prev_groups) {
      fill_u32_kernel<<<agg_launch_grid(ctx, prev_groups, kAggThreads), kAggThreads, 0, ctx->stream>>>(
          (uint32_t*)g->seen->p, prev_groups, 0xffffffffu);
      ctx->launches++;
    }
    g->seen_tracking = true;
  }
  plan.seen = g->seen_tracking ? (uint32_t*)g->seen->p : nullptr;
  for (int32_t i = 0; i < plan.n; i++) plan.a[i].out = (uint64_t*)g->phys[(size_t)i].buf->p;
  if (plan.n == 0) return BGPU_OK;

  // tier by exact group count
  const uint32_t ng = g->n_groups;
  const size_t limit = std::min<size_t>(ctx->smem_optin, 200 * 1024);
  const size_t group_bytes = (size_t)g->total_words * 8;
  const size_t priv = (size_t)ng * group_bytes * kAggThreads + (size_t)ng * 4;
  int grid_cap = ctx->sm_count;
  if (priv <= limit) {
    BGPU_CUDA(ctx, cudaFuncSetAttribute(accumulate_kernel<TIER_PRIVATE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)priv));
    int occ = 1;
    BGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, accumulate_kernel<TIER_PRIVATE>, kAggThreads, priv));
    const int grid = (int)std::min<int64_t>(ceil_div64(n_rows, kAggThreads * kRowsPerIter), (int64_t)grid_cap * std::max(1, occ));
    accumulate_kernel<TIER_PRIVATE><<<grid, kAggThreads, priv, ctx->stream>>>(plan, (const uint32_t*)gids->p, n_rows, ng, 1);
  } else if ((size_t)ng * group_bytes + (size_t)ng * 4 <= limit / 2) {
    int replicas = (int)std::min<size_t>(32, (limit / 2 - (size_t)ng * 4) / ((size_t)ng * group_bytes));
    if (replicas < 1) replicas = 1;
    const size_t smem = (size_t)ng * group_bytes * replicas + (size_t)ng * 4;
    BGPU_CUDA(ctx, cudaFuncSetAttribute(accumulate_kernel<TIER_SHARED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    BGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, accumulate_kernel<TIER_SHARED>, kAggThreads, smem));
    const int grid = (int)std::min<int64_t>(ceil_div64(n_rows, kAggThreads * kRowsPerIter), (int64_t)grid_cap * std::max(1, occ));
    accumulate_kernel<TIER_SHARED><<<grid, kAggThreads, smem, ctx->stream>>>(plan, (const uint32_t*)gids->p, n_rows, ng, replicas);
  } else {
    const int grid = (int)std::min<int64_t>(ceil_div64(n_rows, kAggThreads * kRowsPerIter), (int64_t)grid_cap * 8);
    accumulate_kernel<TIER_GLOBAL><<<grid, kAggThreads, 0, ctx->stream>>>(plan, (const uint32_t*)gids->p, n_rows, ng, 1);
  }
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
}

extern "C" int64_t bgpu_agg_num_groups(bgpu_agg* g) { return g ? (int64_t)g->n_groups : -1; }

#include <cub/device/device_scan.cuh>

static int emit_validity(bgpu_agg* g, int phys_idx, uint32_t n, Column* col) {
  bgpu_ctx* ctx = g->ctx;
  if (!g->seen_tracking || n == 0) return BGPU_OK;
  DevAllocPtr v;
  BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(n, 32) * 4, &v));
  seen_bitmap_kernel<<<agg_launch_grid(ctx, n, kAggThreads), kAggThreads, 0, ctx->stream>>>(
      (const uint32_t*)g->seen->p, n, phys_idx, (uint32_t*)v->p);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  col->validity = (uint8_t*)v->p;
  col->null_count = -1;
  col->owned.push_back(v);
  return BGPU_OK;
}

extern "C" int bgpu_agg_emit(bgpu_agg* g, bgpu_batch** out_batch) {
  if (!g || !out_batch) return BGPU_EINVAL;
  *out_batch = nullptr;
  bgpu_ctx* ctx = g->ctx;
  if (g->emitted) return ctx->fail(BGPU_EINVAL, "aggregate already emitted");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint32_t n = g->n_groups;
  std::unique_ptr<bgpu_batch> out(new bgpu_batch());
  out->ctx = ctx;
  out->n_rows = n;
  const int grid = agg_launch_grid(ctx, std::max<uint32_t>(n, 1), kAggThreads);
  const uint64_t* key0 = (const uint64_t*)g->key0->p;
  const uint64_t* key1 = g->key1 ? (const uint64_t*)g->key1->p : nullptr;
  const uint32_t* kmeta = (const uint32_t*)g->kmeta->p;
  // group key columns
  for (size_t i = 0; i < g->group_cols.size(); i++) {
    const Column& f = g->fields[(size_t)g->group_cols[i]];
    Column c;
    c.name = f.name; c.format = f.format; c.flags = f.flags | ARROW_FLAG_NULLABLE; c.kind = f.kind; c.width = f.width;
    c.length = n; c.null_count = -1;
    DevAllocPtr valid;
    BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(std::max<uint32_t>(n, 1), 32) * 4, &valid));
    c.owned.push_back(valid);
    c.validity = (uint8_t*)valid->p;
    if (f.kind == K_UTF8 || f.kind == K_BINARY) {
      DevAllocPtr lens, offs, data;
      BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &lens));
      BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &offs));
      BGPU_CUDA(ctx, cudaMemsetAsync(offs->p, 0, 4, ctx->stream));
      int32_t total = 0;
      if (n) {
        unpack_utf8_kernel<<<grid, kAggThreads, 0, ctx->stream>>>(key0, key1, kmeta, n, (int)i, g->key_off[i], 0,
                                                                 (int32_t*)lens->p, nullptr, (uint32_t*)valid->p);
        size_t tmp_bytes = 0;
        cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)lens->p, (int32_t*)offs->p + 1, (int)n, ctx->stream);
        DevAllocPtr tmp;
        BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
        BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)lens->p, (int32_t*)offs->p + 1, (int)n, ctx->stream));
        BGPU_CUDA(ctx, cudaMemcpyAsync(&total, (int32_t*)offs->p + n, 4, cudaMemcpyDeviceToHost, ctx->stream));
        BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      }
      BGPU_TRY(dev_alloc(ctx, (size_t)total, &data));
      if (n)
        unpack_utf8_kernel<<<grid, kAggThreads, 0, ctx->stream>>>(key0, key1, kmeta, n, (int)i, g->key_off[i], 1,
                                                                 (int32_t*)offs->p, (uint8_t*)data->p, nullptr);
      ctx->launches += 3;
      c.offsets = (int32_t*)offs->p; c.values = data->p; c.data_bytes = total;
      c.owned.push_back(offs); c.owned.push_back(data);
    } else {
      DevAllocPtr vals;
      const size_t bytes = f.kind == K_BOOL ? (size_t)ceil_div64(std::max<uint32_t>(n, 1), 32) * 4
                                            : (size_t)std::max<uint32_t>(n, 1) * f.width;
      BGPU_TRY(dev_alloc(ctx, bytes, &vals));
      if (n)
        unpack_fixed_kernel<<<grid, kAggThreads, 0, ctx->stream>>>(key0, key1, kmeta, n, (int)i, f.kind, f.width,
                                                                  g->key_off[i], vals->p, (uint32_t*)valid->p);
      ctx->launches++;
      c.values = vals->p;
      c.owned.push_back(vals);
      if (f.kind == K_DICT32) {
        if (g->dicts[i]) {
          c.dict = g->dicts[i];
          c.owned.insert(c.owned.end(), g->dict_owned[i].begin(), g->dict_owned[i].end());
        } else {
          return ctx->fail(BGPU_EUNSUPPORTED, "emit of a dictionary group column before any batch was seen");
        }
      }
    }
    BGPU_CUDA(ctx, cudaGetLastError());
    out->cols.push_back(std::move(c));
  }
  // aggregate columns
  auto acc_column = [&](int phys_idx, const std::string& name, const std::string& format, bool nullable_by_seen,
                        Column* c) -> int {
    const PhysAcc& p = g->phys[(size_t)phys_idx];
    c->name = name; c->format = format; c->flags = ARROW_FLAG_NULLABLE; c->kind = K_FIXED; c->width = p.words * 8;
    c->length = n; c->null_count = 0; c->values = p.buf->p;
    c->owned.push_back(p.buf);  // accumulators are dense arrays: they are the Arrow buffer
    if (nullable_by_seen) BGPU_TRY(emit_validity(g, phys_idx, n, c));
    return BGPU_OK;
  };
  for (const LogicalAgg& a : g->aggs) {
    if (!g->final_mode && g->mode != BGPU_AGG_MODE_SINGLE) {  // Partial: state columns
      if (a.fn == BGPU_AGG_COUNT) {
        Column c; BGPU_TRY(acc_column(a.phys[0], a.name + "[count]", "l", false, &c)); out->cols.push_back(std::move(c));
      } else if (a.fn == BGPU_AGG_SUM) {
        Column c; BGPU_TRY(acc_column(a.phys[0], a.name + "[sum]", a.sum_format, true, &c)); out->cols.push_back(std::move(c));
      } else {
        Column c0; BGPU_TRY(acc_column(a.phys[0], a.name + "[count]", "L", false, &c0)); out->cols.push_back(std::move(c0));
        Column c1; BGPU_TRY(acc_column(a.phys[1], a.name + "[sum]", a.sum_format, true, &c1)); out->cols.push_back(std::move(c1));
      }
      continue;
    }
    if (a.fn == BGPU_AGG_COUNT) {
      Column c; BGPU_TRY(acc_column(a.phys[0], a.name, "l", false, &c)); out->cols.push_back(std::move(c));
    } else if (a.fn == BGPU_AGG_SUM) {
      Column c; BGPU_TRY(acc_column(a.phys[0], a.name, a.result_format, true, &c)); out->cols.push_back(std::move(c));
    } else {  // Avg: evaluate
      Column c;
      c.name = a.name; c.format = a.result_format; c.flags = ARROW_FLAG_NULLABLE; c.kind = K_FIXED;
      c.width = a.decimal ? 16 : 8; c.length = n; c.null_count = -1;
      DevAllocPtr vals, valid;
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<uint32_t>(n, 1) * c.width, &vals));
      BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(std::max<uint32_t>(n, 1), 32) * 4, &valid));
      if (n) {
        avg_evaluate_kernel<<<grid, kAggThreads, 0, ctx->stream>>>(
            (const uint64_t*)g->phys[(size_t)a.phys[1]].buf->p, (const uint64_t*)g->phys[(size_t)a.phys[0]].buf->p, n,
            a.decimal ? 1 : 0, a.scale_up, (uint64_t*)vals->p, (uint32_t*)valid->p, g->d_flags());
        ctx->launches++;
        BGPU_CUDA(ctx, cudaGetLastError());
      }
      c.values = vals->p; c.validity = (uint8_t*)valid->p;
      c.owned.push_back(vals); c.owned.push_back(valid);
      out->cols.push_back(std::move(c));
    }
  }
  uint32_t* h = (uint32_t*)ctx->pinned;
  BGPU_CUDA(ctx, cudaMemcpyAsync(h, g->counters->p, 16, cudaMemcpyDeviceToHost, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (h[3]) return ctx->fail(BGPU_EOVERFLOW, "Arithmetic overflow evaluating Decimal128 AVG");
  g->emitted = true;
  *out_batch = out.release();
  return BGPU_OK;
}

// Kernels of the grouped aggregate (K4 Partial / K5 Final). See aggregate.cu for how the host picks among them.
//
//   fused_private_kernel   ONE pass over the batch for low-cardinality inputs (TPC-H Q1, db-benchmark K=100):
//                          filter -> pack key -> group id from the CTA's shared-memory key cache (global table behind
//                          it) -> evaluate the aggregate arguments -> add into THREAD-PRIVATE accumulators in shared
//                          memory (no atomics, deterministic). Each CTA reduces its partial state into a scratch
//                          slab; commit_kernel folds the slabs into the dense accumulator arrays once the host has
//                          seen that no CTA met a group beyond the private capacity.
//   intern_kernel +        the general two-pass path: pass 1 assigns group ids (idempotent, so a table overflow is
//   accumulate_kernel      handled by growing and re-running), pass 2 adds with shared-memory (replicated) or global
//                          atomics.
#pragma once
#include "agg_device.cuh"

namespace bgpu {

constexpr int kAggThreads = 256;
constexpr uint32_t kNoGroup = 0xffffffffu;  // row dropped by the filter

// ---- two-pass path, pass 1 -----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kAggThreads)
    intern_kernel(const __grid_constant__ KeyPlan kp, const __grid_constant__ GroupTable t,
                  const __grid_constant__ FilterPlan fp, int64_t n_rows, uint32_t* __restrict__ gid_out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    if (fp.n && !eval_filter(fp, r)) {
      gid_out[r] = kNoGroup;
      continue;
    }
    const PackedKey k = pack_key(kp, r, t.flags);
    gid_out[r] = intern(t, k, kp.words);
  }
}

// Rebuild the entry table from the dense keys (after growing).
__global__ void __launch_bounds__(kAggThreads) rehash_kernel(const __grid_constant__ GroupTable t, uint32_t n, int words) {
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) {
    PackedKey k{t.key0[g], words > 1 ? t.key1[g] : 0, t.kmeta[g]};
    const uint64_t h = key_hash(k);
    const unsigned long long tag = (h >> 32) << 32;
    uint32_t i = (uint32_t)h & t.mask;
    while (atomicCAS(&t.entries[i], 0ULL, tag | (unsigned long long)(g + 2u)) != 0ULL) i = (i + 1) & t.mask;
  }
}

// ---- two-pass path, pass 2 -----------------------------------------------------------------------------------------
enum Tier : int { TIER_SHARED = 1, TIER_GLOBAL = 2 };
constexpr int kRowsPerIter = 4;

template <int TIER>
__global__ void __launch_bounds__(kAggThreads)
    accumulate_kernel(const __grid_constant__ ValuePlan vp, const __grid_constant__ AccPlan plan,
                      const uint32_t* __restrict__ gids, int64_t n_rows, uint32_t n_groups, int replicas,
                      int32_t* __restrict__ flags) {
  extern __shared__ __align__(16) uint64_t s_acc[];
  const int W = plan.total_words;
  const int T = kAggThreads;
  uint32_t* s_seen = nullptr;
  if (TIER == TIER_SHARED) {
    for (uint32_t i = threadIdx.x; i < n_groups * W * replicas; i += T) s_acc[i] = 0;
    s_seen = (uint32_t*)(s_acc + (size_t)n_groups * W * replicas);
    if (plan.seen)
      for (uint32_t i = threadIdx.x; i < n_groups; i += T) s_seen[i] = 0;
    __syncthreads();
  }
  const int rep = TIER == TIER_SHARED ? (threadIdx.x % replicas) : 0;
  bool ovf = false;

  const int64_t step = (int64_t)gridDim.x * T * kRowsPerIter;
  for (int64_t base = (int64_t)blockIdx.x * T * kRowsPerIter; base < n_rows; base += step) {
    uint32_t g[kRowsPerIter];
    uint32_t seen_mask[kRowsPerIter];
    Val prev[kRowsPerIter];
    bool prev_ok[kRowsPerIter];
#pragma unroll
    for (int u = 0; u < kRowsPerIter; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      g[u] = r < n_rows ? gids[r] : kNoGroup;
      seen_mask[u] = 0;
      prev[u] = Val{0, 0};
      prev_ok[u] = false;
    }
#pragma unroll 1
    for (int vi = 0; vi < vp.n; vi++) {
      const ValueExpr& e = vp.v[vi];
#pragma unroll
      for (int u = 0; u < kRowsPerIter; u++) {
        if (g[u] == kNoGroup) continue;
        const int64_t r = base + u * T + threadIdx.x;
        bool ok = true;
        const Val v = eval_value(e, r, prev[u], prev_ok[u], &ok, &ovf);
        prev[u] = v;
        prev_ok[u] = ok;
        if (!ok) continue;
#pragma unroll 1
        for (int ai = e.acc_begin; ai < e.acc_end; ai++) {
          const AccK& a = plan.a[ai];
          seen_mask[u] |= 1u << a.seen_bit;
          if (TIER == TIER_SHARED) add_atomic(&s_acc[((size_t)rep * n_groups + g[u]) * W + a.word0], a.op, v);
          else add_atomic(&a.out[(size_t)g[u] * a.words], a.op, v);
        }
      }
    }
    if (plan.seen) {
#pragma unroll
      for (int u = 0; u < kRowsPerIter; u++) {
        if (g[u] == kNoGroup || !seen_mask[u]) continue;
        if (TIER == TIER_GLOBAL) atomicOr(&plan.seen[g[u]], seen_mask[u]);
        else atomicOr(&s_seen[g[u]], seen_mask[u]);
      }
    }
  }
  if (ovf) atomicOr(&flags[2], 1);
  if (TIER == TIER_GLOBAL) return;
  __syncthreads();
  // flush this CTA's replicas: one warp per (group, accumulator)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = T / 32;
  for (uint32_t item = warp; item < n_groups * (uint32_t)plan.n; item += n_warps) {
    const uint32_t grp = item / plan.n;
    const AccK& a = plan.a[item % plan.n];
    unsigned __int128 isum = 0;
    double fsum = 0.0;
    for (int t = lane; t < replicas; t += 32) {
      const uint64_t* p = &s_acc[((size_t)t * n_groups + grp) * W + a.word0];
      if (a.op == OP_ADD_F64) fsum += *(const double*)p;
      else if (a.op == OP_ADD_I128) isum += ((unsigned __int128)p[1] << 64) | p[0];
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
      // partial counts are sums by now
      if (nonzero) add_atomic(&a.out[(size_t)grp * a.words], a.op == OP_COUNT ? OP_ADD_I64 : a.op, v);
    }
  }
  if (plan.seen)
    for (uint32_t i = threadIdx.x; i < n_groups; i += T)
      if (s_seen[i]) atomicOr(&plan.seen[i], s_seen[i]);
}

// ---- fused single-pass path --------------------------------------------------------------------------------------
// Shared memory: acc u64[g_cap * W * T] | seen u32[g_cap] | key cache (slots)
// Scratch slab of CTA b: u64[g_cap * W] partial accumulators, then u32[g_cap] seen masks (padded to 8 bytes).
struct FusedArgs {
  uint64_t* scratch;
  uint32_t g_cap;        // private group capacity: group ids >= g_cap abort the pass
  uint32_t cache_slots;  // power of two
  int32_t rows_per_thread;
  int32_t pad;
};
__host__ __device__ __forceinline__ size_t fused_slab_words(uint32_t g_cap, int W) {
  return (size_t)g_cap * W + ((size_t)g_cap + 1) / 2;
}

template <int T, int R>
__global__ void __launch_bounds__(T)
    fused_private_kernel(const __grid_constant__ KeyPlan kp, const __grid_constant__ GroupTable t,
                         const __grid_constant__ FilterPlan fp, const __grid_constant__ ValuePlan vp,
                         const __grid_constant__ AccPlan plan, const __grid_constant__ FusedArgs fa, int64_t n_rows) {
  extern __shared__ __align__(16) uint64_t smem[];
  const int W = plan.total_words;
  const uint32_t G = fa.g_cap;
  uint64_t* s_acc = smem;
  uint32_t* s_seen = (uint32_t*)(s_acc + (size_t)G * W * T);
  KeyCache cache;
  {
    unsigned char* p = (unsigned char*)(s_seen + ((G + 1) & ~1u));
    cache.k0 = (uint64_t*)p;      p += (size_t)fa.cache_slots * 8;
    cache.k1 = (uint64_t*)p;      p += (size_t)fa.cache_slots * 8;
    cache.state = (uint32_t*)p;   p += (size_t)fa.cache_slots * 4;
    cache.meta = (uint32_t*)p;
    cache.mask = fa.cache_slots - 1;
  }
  __shared__ int s_abort;
  for (uint32_t i = threadIdx.x; i < G * W * T; i += T) s_acc[i] = 0;
  for (uint32_t i = threadIdx.x; i < G; i += T) s_seen[i] = 0;
  for (uint32_t i = threadIdx.x; i < fa.cache_slots; i += T) cache.state[i] = 0;
  if (threadIdx.x == 0) s_abort = 0;
  __syncthreads();

  bool ovf = false;
  uint32_t my_seen_any = 0;
  const int64_t tile = (int64_t)T * R;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < n_rows; base += (int64_t)gridDim.x * tile) {
    uint32_t g[R];
    Val prev[R];
    bool prev_ok[R];
    uint32_t seen_mask[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t r = base + u * T + threadIdx.x;
      g[u] = kNoGroup;
      prev[u] = Val{0, 0};
      prev_ok[u] = false;
      seen_mask[u] = 0;
      if (r < n_rows && (fp.n == 0 || eval_filter(fp, r))) {
        const PackedKey k = pack_key(kp, r, t.flags);
        const uint32_t gid = cached_intern(cache, t, k, kp.words);
        if (gid >= G) s_abort = 1;  // beyond the private capacity (or the global table overflowed)
        else g[u] = gid;
      }
    }
#pragma unroll 1
    for (int vi = 0; vi < vp.n; vi++) {
      const ValueExpr& e = vp.v[vi];
#pragma unroll
      for (int u = 0; u < R; u++) {
        if (g[u] == kNoGroup) continue;
        const int64_t r = base + u * T + threadIdx.x;
        bool ok = true;
        const Val v = eval_value(e, r, prev[u], prev_ok[u], &ok, &ovf);
        prev[u] = v;
        prev_ok[u] = ok;
        if (!ok) continue;
#pragma unroll 1
        for (int ai = e.acc_begin; ai < e.acc_end; ai++) {
          const AccK& a = plan.a[ai];
          if (a.alias >= 0) continue;
          seen_mask[u] |= 1u << a.seen_bit;
          add_private(&s_acc[((size_t)g[u] * W + a.word0) * T + threadIdx.x], T, a.op, v);
        }
      }
    }
    if (plan.seen) {
#pragma unroll
      for (int u = 0; u < R; u++)
        if (g[u] != kNoGroup && seen_mask[u]) atomicOr(&s_seen[g[u]], seen_mask[u]);
    }
    my_seen_any |= 1;
    if (*(volatile int*)&s_abort || *(volatile int*)&t.flags[3]) break;  // some CTA gave up: the pass is void
  }
  if (ovf) atomicOr(&t.flags[2], 1);
  __syncthreads();
  if (s_abort) {
    if (threadIdx.x == 0) atomicOr(&t.flags[3], 1);
    return;
  }
  // reduce the T private copies of every (group, accumulator) into this CTA's scratch slab: one warp per item
  uint64_t* slab = fa.scratch + (size_t)blockIdx.x * fused_slab_words(G, W);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = T / 32;
  for (uint32_t item = warp; item < G * (uint32_t)plan.n; item += n_warps) {
    const uint32_t grp = item / plan.n;
    const AccK& a = plan.a[item % plan.n];
    if (a.alias >= 0) continue;
    unsigned __int128 isum = 0;
    double fsum = 0.0;
    for (int th = lane; th < T; th += 32) {
      const uint64_t* p = &s_acc[((size_t)grp * W + a.word0) * T + th];
      if (a.op == OP_ADD_F64) fsum += *(const double*)p;
      else if (a.op == OP_ADD_I128) isum += ((unsigned __int128)p[T] << 64) | p[0];
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
      uint64_t* o = slab + (size_t)grp * W + a.word0;
      if (a.op == OP_ADD_F64) {
        o[0] = (uint64_t)__double_as_longlong(fsum);
      } else {
        o[0] = (uint64_t)isum;
        if (a.words > 1) o[1] = (uint64_t)(isum >> 64);
      }
    }
  }
  uint32_t* slab_seen = (uint32_t*)(slab + (size_t)G * W);
  for (uint32_t i = threadIdx.x; i < G; i += T) slab_seen[i] = s_seen[i];
}

// Fold the per-CTA slabs into the dense accumulator arrays: one thread per (group, accumulator), slabs in CTA order
// (so Float64 sums are reproducible for a fixed grid).
__global__ void __launch_bounds__(kAggThreads)
    commit_kernel(const __grid_constant__ AccPlan plan, const uint64_t* __restrict__ scratch, uint32_t n_ctas,
                  uint32_t g_cap, uint32_t n_groups) {
  const int W = plan.total_words;
  const size_t slab = fused_slab_words(g_cap, W);
  const uint32_t items = n_groups * (uint32_t)plan.n;
  for (uint32_t item = blockIdx.x * blockDim.x + threadIdx.x; item < items; item += gridDim.x * blockDim.x) {
    const uint32_t grp = item / plan.n;
    const AccK& a = plan.a[item % plan.n];
    const AccK& src = a.alias >= 0 ? plan.a[a.alias] : a;
    unsigned __int128 isum = 0;
    double fsum = 0.0;
    for (uint32_t b = 0; b < n_ctas; b++) {
      const uint64_t* p = scratch + (size_t)b * slab + (size_t)grp * W + src.word0;
      if (src.op == OP_ADD_F64) fsum += __longlong_as_double((long long)p[0]);
      else if (src.op == OP_ADD_I128) isum += ((unsigned __int128)p[1] << 64) | p[0];
      else isum += p[0];
    }
    uint64_t* o = a.out + (size_t)grp * a.words;
    if (a.op == OP_ADD_F64) {
      *(double*)o += fsum;
    } else {
      unsigned __int128 cur = o[0];
      if (a.words > 1) cur |= (unsigned __int128)o[1] << 64;
      cur += isum;
      o[0] = (uint64_t)cur;
      if (a.words > 1) o[1] = (uint64_t)(cur >> 64);
    }
  }
  if (plan.seen) {
    for (uint32_t grp = blockIdx.x * blockDim.x + threadIdx.x; grp < n_groups; grp += gridDim.x * blockDim.x) {
      uint32_t m = 0;
      for (uint32_t b = 0; b < n_ctas; b++) m |= ((const uint32_t*)(scratch + (size_t)b * slab + (size_t)g_cap * W))[grp];
      // aliased accumulators saw whatever their source saw
      for (int i = 0; i < plan.n; i++)
        if (plan.a[i].alias >= 0 && ((m >> plan.a[plan.a[i].alias].seen_bit) & 1)) m |= 1u << plan.a[i].seen_bit;
      if (m) plan.seen[grp] |= m;
    }
  }
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

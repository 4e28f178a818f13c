// Ahead-of-time kernels of the grouped aggregate: the small plan-independent ones (table rebuild, commit of the fused
// path's per-CTA slabs, emit). The hot kernels (fused / intern / accumulate) are specialised per plan at run time:
// agg_jit_prelude.cuh (kernel templates) + agg_codegen.cu (the generated `Plan`).
#pragma once
#include "agg_device.cuh"

namespace bgpu {

constexpr int kAggThreads = 256;

// Re-insert every published slot of the old table with a group id < keep into the (empty) new one after growing;
// group ids are kept (ids >= keep were handed out by the void pass that overflowed and are dropped).
__global__ void __launch_bounds__(kAggThreads)
    rehash_kernel(const uint64_t* __restrict__ old_slots, uint32_t old_capacity, uint32_t keep, const __grid_constant__ GroupTable t) {
  const int W = (int)t.slot_words, smw = W == 2 ? 1 : 2;
  for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < old_capacity; s += gridDim.x * blockDim.x) {
    const uint64_t* o = old_slots + (size_t)s * W;
    const uint64_t sm = o[smw];
    if ((uint32_t)sm < 2u || (uint32_t)sm - 2u >= keep) continue;
    const PackedKey k{o[0], W == 4 ? o[1] : 0ULL, (uint32_t)(sm >> 32)};
    uint32_t i = (uint32_t)key_hash(k) & t.mask;
    while (atomicCAS((unsigned long long*)(slot_ptr(t, i) + smw), 0ULL, (unsigned long long)sm) != 0ULL) i = (i + 1) & t.mask;
    uint64_t* n = slot_ptr(t, i);
    n[0] = k.k0;
    if (W == 4) n[1] = k.k1;
  }
}

// Scratch slab of fused-path CTA b: u64[g_cap * W] partial accumulators, then u32[g_cap] seen masks (padded to 8 B).
__host__ __device__ __forceinline__ size_t fused_slab_words(uint32_t g_cap, int W) {
  return (size_t)g_cap * W + ((size_t)g_cap + 1) / 2;
}

// Fold the per-CTA slabs into the group records: one WARP per (group, accumulator); lanes stride over the slabs and
// the partials are combined by shuffles in a fixed order (Float64 sums are reproducible for a fixed grid).
__device__ __forceinline__ void commit_body(const AccPlan& plan, const uint64_t* __restrict__ scratch, uint32_t n_ctas,
                                            uint32_t g_cap, uint32_t n_groups) {
  const int W = plan.total_words;
  const size_t slab = fused_slab_words(g_cap, W);
  const uint32_t items = n_groups * (uint32_t)plan.n;
  const uint32_t lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t item = warp; item < items; item += n_warps) {
    const uint32_t grp = item / plan.n;
    const AccK& a = plan.a[item % plan.n];
    const AccK& src = a.alias >= 0 ? plan.a[a.alias] : a;
    uint64_t lo = 0, hi = 0;
    double fsum = 0.0;
    for (uint32_t b = lane; b < n_ctas; b += 32) {
      const uint64_t* p = scratch + (size_t)b * slab + (size_t)grp * W + src.word0;
      if (src.op == OP_ADD_F64) {
        fsum += __longlong_as_double((long long)p[0]);
      } else {
        const uint64_t x = p[0];
        lo += x;
        hi += (lo < x) + (src.op == OP_ADD_I128 ? p[1] : 0ULL);
      }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      fsum += __shfl_down_sync(0xffffffffu, fsum, d);
      const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d), ohi = __shfl_down_sync(0xffffffffu, hi, d);
      lo += olo;
      hi += ohi + (lo < olo);
    }
    if (lane == 0) {
      uint64_t* o = plan.acc + (size_t)grp * plan.stride + a.out_off;
      if (a.op == OP_ADD_F64) {
        *(double*)o += fsum;
      } else {
        const uint64_t cur = o[0], nlo = cur + lo;
        o[0] = nlo;
        if (a.words > 1) o[1] += hi + (nlo < cur);
      }
    }
  }
  if (plan.seen) {
    for (uint32_t grp = warp; grp < n_groups; grp += n_warps) {
      uint32_t m = 0;
      for (uint32_t b = lane; b < n_ctas; b += 32) m |= ((const uint32_t*)(scratch + (size_t)b * slab + (size_t)g_cap * W))[grp];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) m |= __shfl_down_sync(0xffffffffu, m, d);
      if (lane == 0) {
        // aliased accumulators saw whatever their source saw
        for (int i = 0; i < plan.n; i++)
          if (plan.a[i].alias >= 0 && ((m >> plan.a[plan.a[i].alias].seen_bit) & 1)) m |= 1u << plan.a[i].seen_bit;
        if (m) plan.seen[grp] |= m;
      }
    }
  }
}

__global__ void __launch_bounds__(kAggThreads)
    commit_kernel(const __grid_constant__ AccPlan plan, const uint64_t* __restrict__ scratch, uint32_t n_ctas,
                  uint32_t g_cap, uint32_t n_groups) {
  commit_body(plan, scratch, n_ctas, g_cap, n_groups);
}

// The same for a launch whose outcome the host has not looked at (streamed input, ingest.cu): the device decides. The
// slabs are folded only when the fused launch neither overflowed the table (counters[1]) nor met a group beyond its
// capacity (counters[4]) - both flags are sticky until the host clears them, so every later launch is skipped as well
// and the host replays the skipped chunks in order. status = {committed, n_groups, counters[1..6]}.
__global__ void __launch_bounds__(kAggThreads)
    commit_cond_kernel(const __grid_constant__ AccPlan plan, const uint64_t* __restrict__ scratch, uint32_t n_ctas,
                       uint32_t g_cap, const uint32_t* __restrict__ counters, uint32_t* __restrict__ status) {
  const uint32_t n_groups = counters[0];
  const bool ok = counters[1] == 0 && counters[4] == 0 && n_groups <= g_cap;
  if (blockIdx.x == 0 && threadIdx.x < 8)
    status[threadIdx.x] = threadIdx.x == 0 ? (ok ? 1u : 0u) : threadIdx.x == 1 ? n_groups : counters[threadIdx.x - 1];
  if (ok) commit_body(plan, scratch, n_ctas, g_cap, n_groups);
}

// ---- emit kernels -----------------------------------------------------------------------------------------------
// Dense key arrays (index = group id) from the table slots.
__global__ void __launch_bounds__(kAggThreads)
    densify_keys_kernel(const __grid_constant__ GroupTable t, uint64_t* __restrict__ key0, uint64_t* __restrict__ key1,
                        uint32_t* __restrict__ kmeta) {
  const int W = (int)t.slot_words, smw = W == 2 ? 1 : 2;
  for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s <= t.mask; s += gridDim.x * blockDim.x) {
    const uint64_t* o = t.slots + (size_t)s * W;
    const uint64_t sm = o[smw];
    if ((uint32_t)sm < 2u) continue;
    const uint32_t gid = (uint32_t)sm - 2u;
    if (gid >= t.max_groups) continue;
    key0[gid] = o[0];
    if (key1) key1[gid] = W == 4 ? o[1] : 0ULL;
    kmeta[gid] = (uint32_t)(sm >> 32);
  }
}

// One accumulator of every group record -> a dense Arrow values buffer (words = 1 or 2).
__global__ void __launch_bounds__(kAggThreads)
    extract_acc_kernel(const uint64_t* __restrict__ acc, uint32_t stride, int off, int words, uint32_t n, uint64_t* __restrict__ out) {
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) {
    const uint64_t* p = acc + (size_t)g * stride + off;
    out[(size_t)g * words] = p[0];
    if (words > 1) out[(size_t)g * words + 1] = p[1];
  }
}

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

// set bits of the cardinality sketch's bitmap (run_two_pass)
__global__ void __launch_bounds__(kAggThreads) popcount_kernel(const uint32_t* __restrict__ w, uint32_t n, unsigned long long* out) {
  unsigned long long c = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) c += __popc(w[i]);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) c += __shfl_down_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// OR of all group key metas: bit c set = some group has a null in key column c (emit drops all-valid bitmaps)
__global__ void __launch_bounds__(kAggThreads) or_kmeta_kernel(const uint32_t* __restrict__ kmeta, uint32_t n, uint32_t* __restrict__ out) {
  uint32_t v = 0;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) v |= kmeta[g];
  v = __reduce_or_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0 && v) atomicOr(out, v);
}

__global__ void __launch_bounds__(kAggThreads) fill_u32_kernel(uint32_t* p, uint32_t n, uint32_t v) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = v;
}

// AvgGroupsAccumulator::evaluate from the group records. Float64: sum / count. Decimal128: (sum * 10^scale_up) / count,
// truncating; an overflowing scaled multiply is an error in the reference -> flags[2].
__global__ void __launch_bounds__(kAggThreads)
    avg_evaluate_kernel(const uint64_t* __restrict__ acc, uint32_t stride, int sum_off, int count_off, uint32_t n, int is_decimal,
                        int scale_up, uint64_t* __restrict__ out, uint32_t* __restrict__ out_valid, int32_t* flags) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false;
    if (g < n) {
      const uint64_t* rec = acc + (size_t)g * stride;
      const uint64_t c = rec[count_off];
      valid = c != 0;
      if (is_decimal) {
        __int128 s = (__int128)(((unsigned __int128)rec[sum_off + 1] << 64) | rec[sum_off]);
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
        const double s = __longlong_as_double((long long)rec[sum_off]);
        out[g] = (uint64_t)__double_as_longlong(valid ? s / (double)c : 0.0);
      }
    }
    const uint32_t vb = __ballot_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
  }
}
}  // namespace bgpu

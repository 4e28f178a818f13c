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
  const bool ok = counters[1] == 0 && counters[2] == 0 && counters[3] == 0 && counters[4] == 0 && n_groups <= g_cap;
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

// ---- emit in one launch ---------------------------------------------------------------------------------------------
// Every output column of the state / result batch from the dense key arrays and the group records: key columns
// (fixed-width, bool, dictionary indices, utf8 lengths), accumulator columns (1 or 2 words per group), AVG evaluation and
// all validity bitmaps (ballots: thread = group id). SINGLE = one CTA covers all groups and also scans the utf8 lengths
// into offsets and writes the string bytes; otherwise the host runs a device scan + unpack_utf8_kernel for utf8 keys.
constexpr int kEmitThreads = 1024;
constexpr int kMaxEmitAcc = 40;
struct EmitKeyCol {
  void* values;       // fixed / bool / dictionary indices
  int32_t* offsets;   // utf8: n + 1 entries
  int32_t* lens;      // utf8 (not SINGLE): n entries
  uint8_t* data;      // utf8 bytes
  uint32_t* valid;    // always present; the host drops it when no group has a null in this column
  int32_t kind, width, byte_off, pad;
};
struct EmitAccCol {
  uint64_t* out;
  uint32_t* valid;    // nullptr: never null
  int32_t type;       // 0: copy `words` words at `off`; 1: AVG Float64 (sum at off, count at count_off); 2: AVG Decimal128
  int32_t words, off, count_off, scale_up, seen_bit;
};
struct EmitArgs {
  EmitKeyCol k[kMaxGroupCols];
  EmitAccCol a[kMaxEmitAcc];
  const uint64_t* key0;
  const uint64_t* key1;
  const uint32_t* kmeta;
  const uint64_t* acc;
  const uint32_t* seen;
  int32_t* flags;          // the aggregate's counters + 1
  uint32_t* key_null_or;   // OR of all key metas (bit c = some group has a null in key column c)
  uint32_t stride;
  int32_t n_key, n_acc;
};

template <bool SINGLE>
__global__ void __launch_bounds__(kEmitThreads) emit_all_kernel(const __grid_constant__ EmitArgs a, uint32_t n) {
  const uint32_t n_pad = (n + 31) & ~31u;
  const bool lane0 = (threadIdx.x & 31) == 0;
  uint32_t null_or = 0;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    const bool in = g < n;
    uint64_t k0 = 0, k1 = 0;
    uint32_t meta = 0;
    if (in) {
      k0 = a.key0[g];
      k1 = a.key1 ? a.key1[g] : 0;
      meta = a.kmeta[g];
      null_or |= meta;
    }
    for (int c = 0; c < a.n_key; c++) {
      const EmitKeyCol& kc = a.k[c];
      const bool valid = in && !((meta >> c) & 1);
      const uint64_t word = kc.byte_off < 8 ? (k0 >> (8 * kc.byte_off)) : (k1 >> (8 * (kc.byte_off - 8)));
      if (kc.kind == K_UTF8 || kc.kind == K_BINARY) {
        if (in) (SINGLE ? kc.offsets + 1 : kc.lens)[g] = valid ? (int32_t)(word & 0xff) : 0;
      } else if (kc.kind == K_BOOL) {
        const uint32_t bits = __ballot_sync(0xffffffffu, in && (word & 1));
        if (lane0) ((uint32_t*)kc.values)[g >> 5] = bits;
      } else if (in) {
        if (kc.width == 16) ((ulonglong2*)kc.values)[g] = make_ulonglong2(k0, k1);
        else if (kc.width == 8) ((uint64_t*)kc.values)[g] = word;
        else if (kc.width == 4) ((uint32_t*)kc.values)[g] = (uint32_t)word;
        else if (kc.width == 2) ((uint16_t*)kc.values)[g] = (uint16_t)word;
        else ((uint8_t*)kc.values)[g] = (uint8_t)word;
      }
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if (lane0) kc.valid[g >> 5] = vb;
    }
    const uint64_t* rec = a.acc + (size_t)(in ? g : 0) * a.stride;
    const uint32_t seen = (a.seen && in) ? a.seen[g] : 0xffffffffu;
    for (int j = 0; j < a.n_acc; j++) {
      const EmitAccCol& ac = a.a[j];
      bool valid = in;
      if (ac.type == 0) {
        if (in) {
          ac.out[(size_t)g * ac.words] = rec[ac.off];
          if (ac.words > 1) ac.out[(size_t)g * ac.words + 1] = rec[ac.off + 1];
        }
        valid = in && ((seen >> ac.seen_bit) & 1);
      } else {
        const uint64_t cnt = in ? rec[ac.count_off] : 0;
        valid = in && cnt != 0;
        if (in && ac.type == 2) {
          __int128 s = (__int128)(((unsigned __int128)rec[ac.off + 1] << 64) | rec[ac.off]);
          __int128 q = 0;
          if (valid) {
            bool ovf = false;
            for (int i = 0; i < ac.scale_up; i++) {
              const __int128 lim = (((__int128)1 << 126) - 1 + ((__int128)1 << 126)) / 10;  // i128::MAX / 10
              if (s > lim || s < -lim) ovf = true;
              s *= 10;
            }
            if (ovf) atomicOr(&a.flags[2], 1);
            q = s / (__int128)cnt;
          }
          ac.out[2 * (size_t)g] = (uint64_t)q;
          ac.out[2 * (size_t)g + 1] = (uint64_t)((unsigned __int128)q >> 64);
        } else if (in) {
          const double s = __longlong_as_double((long long)rec[ac.off]);
          ac.out[g] = (uint64_t)__double_as_longlong(valid ? s / (double)cnt : 0.0);
        }
      }
      if (ac.valid) {
        const uint32_t vb = __ballot_sync(0xffffffffu, valid);
        if (lane0) ac.valid[g >> 5] = vb;
      }
    }
  }
  null_or = __reduce_or_sync(0xffffffffu, null_or);
  if (lane0 && null_or) atomicOr(a.key_null_or, null_or);
  if (!SINGLE) return;
  // one CTA: lengths (parked in offsets[1..n]) -> offsets, then the bytes
  __shared__ int32_t s_warp[kEmitThreads / 32];
  __shared__ int32_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int c = 0; c < a.n_key; c++) {
    const EmitKeyCol& kc = a.k[c];
    if (kc.kind != K_UTF8 && kc.kind != K_BINARY) continue;
    __syncthreads();
    if (threadIdx.x == 0) { s_carry = 0; kc.offsets[0] = 0; }
    __syncthreads();
    for (uint32_t g0 = 0; g0 < n; g0 += kEmitThreads) {
      const uint32_t g = g0 + threadIdx.x;
      const int32_t len = g < n ? kc.offsets[g + 1] : 0;
      int32_t x = len;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
      }
      if (lane == 31) s_warp[warp] = x;
      __syncthreads();
      int32_t pre = s_carry;
      for (int w = 0; w < warp; w++) pre += s_warp[w];
      const int32_t end = pre + x;
      if (g < n) {
        kc.offsets[g + 1] = end;
        const uint64_t k0 = a.key0[g], k1 = a.key1 ? a.key1[g] : 0;
        for (int i = 0; i < len; i++) {
          const int off = kc.byte_off + 1 + i;
          kc.data[end - len + i] = (uint8_t)((off < 8 ? (k0 >> (8 * off)) : (k1 >> (8 * (off - 8)))) & 0xff);
        }
      }
      __syncthreads();
      if (threadIdx.x == kEmitThreads - 1) s_carry = end;
      __syncthreads();
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

__global__ void __launch_bounds__(kAggThreads) fill_u64_kernel(uint64_t* p, int64_t n, uint64_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
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

// K1-K3: create_hashes, partition ids + histogram, scan, stable scatter (take-by-partition).
//
// Replaces DataFusion's BatchPartitioner::partition hash arm as driven by ShuffleWriterExec
// (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-391):
//   hash = create_hashes(keys, seeds 0,0,0,0); partition = hash % n; per partition, rows in input order.
//
// Layout: the batch is cut into G contiguous row chunks, one per persistent CTA (G = resident CTAs).
//   pass 1 (count):   each CTA histograms its chunk's partition ids in shared memory -> hist[p][chunk]
//   scan:             exclusive scan of hist in (partition, chunk) order -> base[p][chunk], offsets[p]
//   pass 2 (scatter): each CTA walks its chunk in tiles of kTile rows; a tile is ranked stably by partition
//                     with warp match + per-warp counters, every column is staged through shared memory in
//                     partition order and written out in runs (coalesced), validity bits with atomicOr.
// HBM traffic per row: key columns twice (pass 1 + pass 2, which scatters them anyway), other columns once,
// every output byte once: 88 B/row for the Int64-key + 4x8 B payload microbench vs 80 B/row algorithmic.
#include <cub/device/device_scan.cuh>

#include "hash.cuh"
#include "internal.h"

namespace bgpu {

#ifndef BGPU_PART_THREADS
#define BGPU_PART_THREADS 256
#define BGPU_PART_ITEMS 16
#define BGPU_PART_MIN_CTAS 3
#endif
constexpr int kThreads = BGPU_PART_THREADS;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = BGPU_PART_ITEMS;
constexpr int kTile = kThreads * kItems;  // 4096 rows
constexpr uint32_t kMaxPartitions = 2048;

// ---- K1: create_hashes -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) hash_kernel(const __grid_constant__ KeySet ks, int64_t n_rows,
                                                        uint64_t* __restrict__ out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x)
    out[r] = row_hash(ks, r);
}

// ---- K2: partition ids + per-chunk histogram ----------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
    count_kernel(const __grid_constant__ KeySet ks, const __grid_constant__ Modulo mod, int64_t n_rows,
                 int64_t chunk_rows, uint32_t* __restrict__ hist /* [n][gridDim.x] */,
                 uint64_t* __restrict__ out_hashes) {
  extern __shared__ uint32_t s_hist[];
  const uint32_t n = mod.n;
  for (uint32_t p = threadIdx.x; p < n; p += kThreads) s_hist[p] = 0;
  __syncthreads();
  const int64_t begin = (int64_t)blockIdx.x * chunk_rows;
  int64_t end = begin + chunk_rows;
  if (end > n_rows) end = n_rows;
  for (int64_t r = begin + threadIdx.x; r < end; r += kThreads) {
    const uint64_t h = row_hash(ks, r);
    if (out_hashes) out_hashes[r] = h;
    atomicAdd(&s_hist[partition_of(h, mod)], 1u);
  }
  __syncthreads();
  for (uint32_t p = threadIdx.x; p < n; p += kThreads) hist[(size_t)p * gridDim.x + blockIdx.x] = s_hist[p];
}

// ---- scan of hist[p][chunk] (single CTA; n*G <= ~1.2M entries) ---------------------------------------------
__global__ void __launch_bounds__(1024) scan_kernel(const uint32_t* __restrict__ hist, uint64_t* __restrict__ base,
                                                    uint64_t* __restrict__ offsets, uint32_t n, uint32_t g,
                                                    int64_t n_rows) {
  __shared__ uint64_t s_warp[32];
  __shared__ uint64_t s_carry;
  const int64_t total = (int64_t)n * g;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int64_t i0 = 0; i0 < total; i0 += 1024) {
    const int64_t i = i0 + threadIdx.x;
    const uint64_t v = i < total ? hist[i] : 0;
    uint64_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint64_t w = s_warp[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint64_t y = __shfl_up_sync(0xffffffffu, w, d);
        if (lane >= d) w += y;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint64_t carry = s_carry;
    const uint64_t excl = carry + (warp ? s_warp[warp - 1] : 0) + x - v;
    if (i < total) {
      base[i] = excl;
      if (i % g == 0) offsets[i / g] = excl;
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + s_warp[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[n] = (uint64_t)n_rows;
}

// ---- K3: stable scatter ---------------------------------------------------------------------------------------
constexpr int kMaxCols = 32;
struct ScatterCol {
  const void* in;           // fixed values / dict keys / bool bits
  void* out;
  const uint8_t* in_valid;  // nullptr = no nulls
  uint32_t* out_valid;      // zero-initialised, 32-bit words
  const int32_t* in_offsets;  // utf8/binary: the scattered "value" is (start, length)
  int32_t* out_len;         // utf8: lengths land in out_len[d] (scanned into offsets afterwards)
  int32_t* out_src;         // utf8: source byte start of output row d
  int32_t kind;
  int32_t width;
};
struct ScatterArgs {
  ScatterCol cols[kMaxCols];
  int32_t n_cols;
  int32_t stage_width;  // bytes per staged element: widest scattered value, multiple of 4, <= 16
};

// peers of this lane's partition id within the warp
__device__ __forceinline__ uint32_t match_peers(uint32_t pid) { return __match_any_sync(0xffffffffu, pid); }

// Stage one column's values of this tile into shared memory in partition order, then write them out in runs.
template <typename T, typename Load>
__device__ __forceinline__ void scatter_staged(Load load, T* __restrict__ out, T* stage, const uint16_t* pid_sorted,
                                               const int64_t* gdelta, int64_t tile_base, int tile_n,
                                               const uint32_t (&pos)[kItems]) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kBatch = sizeof(T) >= 16 ? 4 : (sizeof(T) >= 8 ? 8 : 16);  // loads in flight per thread
#pragma unroll
  for (int b0 = 0; b0 < kItems; b0 += kBatch) {
    T v[kBatch];
#pragma unroll
    for (int k = 0; k < kBatch; k++) {
      const int local = w * (32 * kItems) + (b0 + k) * 32 + lane;
      if (local < tile_n) v[k] = load(tile_base + local);
    }
#pragma unroll
    for (int k = 0; k < kBatch; k++) {
      const int local = w * (32 * kItems) + (b0 + k) * 32 + lane;
      if (local < tile_n) stage[pos[b0 + k]] = v[k];
    }
  }
  __syncthreads();
#pragma unroll 4
  for (int j = threadIdx.x; j < tile_n; j += kThreads) out[j + gdelta[pid_sorted[j]]] = stage[j];
  __syncthreads();
}

// Scatter a bitmap (validity or boolean values): set bit d of the zero-initialised output for every set input bit.
__device__ __forceinline__ void scatter_bits(const uint8_t* __restrict__ in, uint32_t* __restrict__ out,
                                             const int64_t* gdelta, const uint32_t (&pid)[kItems], int64_t tile_base,
                                             int tile_n, const uint32_t (&pos)[kItems]) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int it = 0; it < kItems; it++) {
    const int local = w * (32 * kItems) + it * 32 + lane;
    if (local < tile_n && bit_is_set(in, tile_base + local)) {
      const int64_t d = (int64_t)pos[it] + gdelta[pid[it]];
      atomicOr(&out[d >> 5], 1u << (d & 31));
    }
  }
}

__global__ void __launch_bounds__(kThreads, BGPU_PART_MIN_CTAS)
    scatter_kernel(const __grid_constant__ KeySet ks, const __grid_constant__ Modulo mod,
                   const __grid_constant__ ScatterArgs args, int64_t n_rows, int64_t chunk_rows,
                   const uint64_t* __restrict__ base /* [n][gridDim.x] */) {
  extern __shared__ __align__(16) unsigned char smem[];
  const uint32_t n = mod.n;
  // shared layout: stage (16 B aligned) | gdelta i64[n] | chunk_base u64[n] | warp_cnt u32[kWarps*n] |
  //                tile_off u32[n] | tile_cnt u32[n] | pid_sorted u16[kTile]
  unsigned char* sp = smem;
  void* stage = sp;                       sp += (size_t)kTile * args.stage_width;
  int64_t* gdelta = (int64_t*)sp;         sp += (size_t)n * 8;
  uint64_t* chunk_base = (uint64_t*)sp;   sp += (size_t)n * 8;
  uint32_t* warp_cnt = (uint32_t*)sp;     sp += (size_t)kWarps * n * 4;
  uint32_t* tile_off = (uint32_t*)sp;     sp += (size_t)n * 4;
  uint32_t* tile_cnt = (uint32_t*)sp;     sp += (size_t)n * 4;
  uint16_t* pid_sorted = (uint16_t*)sp;
  __shared__ uint32_t s_scan[kWarps];

  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1;
  const int64_t begin = (int64_t)blockIdx.x * chunk_rows;
  int64_t end = begin + chunk_rows;
  if (end > n_rows) end = n_rows;
  for (uint32_t p = threadIdx.x; p < n; p += kThreads) chunk_base[p] = base[(size_t)p * gridDim.x + blockIdx.x];

  for (int64_t tile_base = begin; tile_base < end; tile_base += kTile) {
    const int tile_n = (int)((end - tile_base) < kTile ? (end - tile_base) : kTile);
    for (uint32_t i = threadIdx.x; i < kWarps * n; i += kThreads) warp_cnt[i] = 0;
    __syncthreads();
    // 1. partition id and stable rank within this warp's segment
    uint32_t pid[kItems], pos[kItems];
    uint32_t* wc = warp_cnt + (size_t)w * n;
#pragma unroll
    for (int it = 0; it < kItems; it++) {
      const int local = w * (32 * kItems) + it * 32 + lane;
      const bool valid = local < tile_n;
      pid[it] = valid ? partition_of(row_hash(ks, tile_base + local), mod) : 0xffffffffu;
    }
#pragma unroll
    for (int it = 0; it < kItems; it++) {
      const uint32_t peers = match_peers(pid[it]);
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if (lane == leader && pid[it] != 0xffffffffu) {
        old = wc[pid[it]];
        wc[pid[it]] = old + __popc(peers);
      }
      old = __shfl_sync(0xffffffffu, old, leader);
      pos[it] = old + __popc(peers & lt_mask);
      __syncwarp();
    }
    __syncthreads();
    // 2. per partition: exclusive scan over warps, tile totals, then exclusive scan over partitions
    for (uint32_t p = threadIdx.x; p < n; p += kThreads) {
      uint32_t run = 0;
#pragma unroll
      for (int ww = 0; ww < kWarps; ww++) {
        const uint32_t c = warp_cnt[(size_t)ww * n + p];
        warp_cnt[(size_t)ww * n + p] = run;
        run += c;
      }
      tile_cnt[p] = run;
    }
    __syncthreads();
    {  // block exclusive scan of tile_cnt[0..n) into tile_off; each thread owns a contiguous slice
      const uint32_t per = (n + kThreads - 1) / kThreads;
      const uint32_t p0 = threadIdx.x * per;
      uint32_t sum = 0;
      for (uint32_t k = 0; k < per; k++)
        if (p0 + k < n) sum += tile_cnt[p0 + k];
      uint32_t x = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
      }
      if (lane == 31) s_scan[w] = x;
      __syncthreads();
      uint32_t wpre = 0;
      for (int ww = 0; ww < w; ww++) wpre += s_scan[ww];
      uint32_t run = wpre + x - sum;
      for (uint32_t k = 0; k < per; k++)
        if (p0 + k < n) {
          const uint32_t p = p0 + k;
          tile_off[p] = run;
          gdelta[p] = (int64_t)chunk_base[p] - (int64_t)run;
          chunk_base[p] += tile_cnt[p];
          run += tile_cnt[p];
        }
    }
    __syncthreads();
    // 3. position of every row in the partition-ordered tile
#pragma unroll
    for (int it = 0; it < kItems; it++) {
      if (pid[it] != 0xffffffffu) {
        pos[it] += tile_off[pid[it]] + wc[pid[it]];
        pid_sorted[pos[it]] = (uint16_t)pid[it];
      }
    }
    __syncthreads();
    // 4. columns
    for (int c = 0; c < args.n_cols; c++) {
      const ScatterCol& col = args.cols[c];
      if (col.in_valid) scatter_bits(col.in_valid, col.out_valid, gdelta, pid, tile_base, tile_n, pos);
      if (col.kind == K_BOOL) {
        scatter_bits((const uint8_t*)col.in, (uint32_t*)col.out, gdelta, pid, tile_base, tile_n, pos);
      } else if (col.kind == K_UTF8 || col.kind == K_BINARY) {
        const int32_t* offs = col.in_offsets;
        scatter_staged<int32_t>([=](int64_t r) { return offs[r + 1] - offs[r]; }, col.out_len, (int32_t*)stage,
                                pid_sorted, gdelta, tile_base, tile_n, pos);
        scatter_staged<int32_t>([=](int64_t r) { return offs[r]; }, col.out_src, (int32_t*)stage, pid_sorted, gdelta,
                                tile_base, tile_n, pos);
      } else {
        switch (col.width) {
          case 8: {
            const uint64_t* in = (const uint64_t*)col.in;
            scatter_staged<uint64_t>([=](int64_t r) { return in[r]; }, (uint64_t*)col.out, (uint64_t*)stage,
                                     pid_sorted, gdelta, tile_base, tile_n, pos);
          } break;
          case 16: {
            const uint4* in = (const uint4*)col.in;
            scatter_staged<uint4>([=](int64_t r) { return in[r]; }, (uint4*)col.out, (uint4*)stage, pid_sorted,
                                  gdelta, tile_base, tile_n, pos);
          } break;
          case 4: {
            const uint32_t* in = (const uint32_t*)col.in;
            scatter_staged<uint32_t>([=](int64_t r) { return in[r]; }, (uint32_t*)col.out, (uint32_t*)stage,
                                     pid_sorted, gdelta, tile_base, tile_n, pos);
          } break;
          case 2: {
            const uint16_t* in = (const uint16_t*)col.in;
            scatter_staged<uint16_t>([=](int64_t r) { return in[r]; }, (uint16_t*)col.out, (uint16_t*)stage,
                                     pid_sorted, gdelta, tile_base, tile_n, pos);
          } break;
          default: {
            const uint8_t* in = (const uint8_t*)col.in;
            scatter_staged<uint8_t>([=](int64_t r) { return in[r]; }, (uint8_t*)col.out, (uint8_t*)stage,
                                    pid_sorted, gdelta, tile_base, tile_n, pos);
          } break;
        }
      }
    }
  }
}

// utf8/binary payload: copy row bytes once output offsets are known (one thread per output row)
__global__ void __launch_bounds__(kThreads)
    copy_bytes_kernel(const uint8_t* __restrict__ in_data, const int32_t* __restrict__ src_start,
                      const int32_t* __restrict__ out_offsets, uint8_t* __restrict__ out_data, int64_t n_rows) {
  for (int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; d < n_rows; d += (int64_t)gridDim.x * blockDim.x) {
    const int32_t b = out_offsets[d], e = out_offsets[d + 1];
    const uint8_t* src = in_data + src_start[d];
    for (int32_t i = 0; i < e - b; i++) out_data[b + i] = src[i];
  }
}

static Modulo make_modulo(uint32_t n) {
  Modulo m{};
  m.n = n;
  m.pow2 = (n & (n - 1)) == 0;
  m.mask = n - 1;
  m.small = n < 65536;
  m.magic = n > 1 ? (UINT64_MAX / n + 1) : 0;
  m.two32_mod = (uint32_t)((1ULL << 32) % n);
  return m;
}

// Build the kernel view of the key columns; dictionary keys get their dictionary hashed once (hash_dictionary).
static int make_keyset(bgpu_ctx* ctx, const bgpu_batch* b, const int32_t* key_cols, int32_t n_keys, KeySet* ks,
                       std::vector<DevAllocPtr>* temps) {
  if (n_keys < 1 || n_keys > kMaxKeys) return ctx->fail(BGPU_EUNSUPPORTED, "1..8 hash key columns supported");
  ks->n = n_keys;
  for (int32_t i = 0; i < n_keys; i++) {
    const int32_t ci = key_cols[i];
    if (ci < 0 || ci >= (int32_t)b->cols.size()) return ctx->fail(BGPU_EINVAL, "key column index out of range");
    const Column& c = b->cols[(size_t)ci];
    KCol& k = ks->cols[i];
    k = KCol{c.values, c.offsets, c.validity, nullptr, c.kind, c.width};
    if (c.kind == K_DICT32) {
      const Column& d = *c.dict;
      if (d.kind == K_DICT32) return ctx->fail(BGPU_EUNSUPPORTED, "nested dictionary");
      DevAllocPtr dh;
      BGPU_TRY(dev_alloc(ctx, (size_t)d.length * 8, &dh));
      temps->push_back(dh);
      if (d.length) {
        KeySet dks{};
        dks.n = 1;
        dks.cols[0] = KCol{d.values, d.offsets, d.validity, nullptr, d.kind, d.width};
        const int grid = (int)std::min<int64_t>(ceil_div64(d.length, kThreads), ctx->sm_count * 8);
        hash_kernel<<<grid, kThreads, 0, ctx->stream>>>(dks, d.length, (uint64_t*)dh->p);
        ctx->launches++;
        BGPU_CUDA(ctx, cudaGetLastError());
      }
      k.dict_hashes = (const uint64_t*)dh->p;
    }
  }
  ks->fast64 = (n_keys == 1 && ks->cols[0].kind == K_FIXED && ks->cols[0].width == 8 && !ks->cols[0].validity);
  return BGPU_OK;
}
}  // namespace bgpu

using namespace bgpu;

extern "C" int bgpu_create_hashes(bgpu_ctx* ctx, const bgpu_batch* batch, const int32_t* key_cols, int32_t n_keys,
                                  uint64_t* out_hashes_device) {
  if (!ctx || !batch || !key_cols || !out_hashes_device) return BGPU_EINVAL;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  KeySet ks{};
  std::vector<DevAllocPtr> temps;
  BGPU_TRY(make_keyset(ctx, batch, key_cols, n_keys, &ks, &temps));
  if (batch->n_rows == 0) return BGPU_OK;
  const int grid = (int)std::min<int64_t>(ceil_div64(batch->n_rows, kThreads), ctx->sm_count * 8);
  hash_kernel<<<grid, kThreads, 0, ctx->stream>>>(ks, batch->n_rows, out_hashes_device);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
}

static size_t scatter_smem_bytes(uint32_t n, int stage_width) {
  return (size_t)kTile * stage_width + (size_t)n * 8 * 2 + (size_t)kWarps * n * 4 + (size_t)n * 4 * 2 + (size_t)kTile * 2;
}

extern "C" int bgpu_hash_partition(bgpu_ctx* ctx, const bgpu_batch* in, const int32_t* key_cols, int32_t n_keys,
                                   uint32_t n_partitions, bgpu_batch** out_partitioned, uint64_t* out_offsets,
                                   uint64_t* out_hashes_device) {
  if (!ctx || !in || !key_cols || !out_partitioned || !out_offsets) return BGPU_EINVAL;
  *out_partitioned = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (n_partitions == 0) return ctx->fail(BGPU_EINVAL, "n_partitions must be > 0");
  if (n_partitions > kMaxPartitions)
    return ctx->fail(BGPU_EUNSUPPORTED, "n_partitions > " + std::to_string(kMaxPartitions));
  if (in->cols.size() > (size_t)kMaxCols) return ctx->fail(BGPU_EUNSUPPORTED, "more than 32 columns");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t n_rows = in->n_rows;
  const uint32_t n = n_partitions;

  std::unique_ptr<bgpu_batch> out(new bgpu_batch(ctx));
  out->n_rows = n_rows;
  out->cols.resize(in->cols.size());

  // output columns + scatter descriptors
  ScatterArgs args{};
  args.n_cols = (int32_t)in->cols.size();
  int stage_width = 4;
  struct Utf8Post { size_t col; DevAllocPtr len, src; };
  std::vector<Utf8Post> utf8;
  for (size_t i = 0; i < in->cols.size(); i++) {
    const Column& c = in->cols[i];
    Column& o = out->cols[i];
    o.name = c.name; o.format = c.format; o.flags = c.flags; o.kind = c.kind; o.width = c.width;
    o.length = n_rows; o.null_count = c.null_count; o.dict = c.dict;
    ScatterCol& sc = args.cols[i];
    sc = ScatterCol{};
    sc.kind = c.kind; sc.width = c.width;
    if (c.validity) {
      DevAllocPtr v;
      const size_t vb = (size_t)ceil_div64(n_rows, 32) * 4;
      BGPU_TRY(dev_alloc(ctx, vb, &v));
      BGPU_CUDA(ctx, cudaMemsetAsync(v->p, 0, v->bytes, ctx->stream));
      o.owned.push_back(v); o.validity = (uint8_t*)v->p;
      sc.in_valid = c.validity; sc.out_valid = (uint32_t*)v->p;
    }
    DevAllocPtr a;
    if (c.kind == K_UTF8 || c.kind == K_BINARY) {
      BGPU_TRY(dev_alloc(ctx, (size_t)(n_rows + 1) * 4, &a));
      o.owned.push_back(a); o.offsets = (int32_t*)a->p;
      Utf8Post up; up.col = i;
      BGPU_TRY(dev_alloc(ctx, (size_t)(n_rows + 1) * 4, &up.len));
      BGPU_TRY(dev_alloc(ctx, (size_t)(n_rows + 1) * 4, &up.src));
      sc.in_offsets = c.offsets; sc.out_len = (int32_t*)up.len->p; sc.out_src = (int32_t*)up.src->p;
      utf8.push_back(up);
    } else if (c.kind == K_BOOL) {
      BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(n_rows, 32) * 4, &a));
      BGPU_CUDA(ctx, cudaMemsetAsync(a->p, 0, a->bytes, ctx->stream));
      o.owned.push_back(a); o.values = a->p;
      sc.in = c.values; sc.out = a->p;
    } else {
      if (c.width != 1 && c.width != 2 && c.width != 4 && c.width != 8 && c.width != 16)
        return ctx->fail(BGPU_EUNSUPPORTED, "unsupported fixed width");
      BGPU_TRY(dev_alloc(ctx, (size_t)n_rows * c.width, &a));
      o.owned.push_back(a); o.values = a->p;
      sc.in = c.values; sc.out = a->p;
      if (c.width > stage_width) stage_width = c.width;
      if (c.kind == K_DICT32) o.owned.insert(o.owned.end(), c.owned.begin(), c.owned.end());  // shares the dictionary
    }
  }
  args.stage_width = stage_width;

  if (n_rows == 0) {
    for (uint32_t p = 0; p <= n; p++) out_offsets[p] = 0;
    for (auto& up : utf8) BGPU_CUDA(ctx, cudaMemsetAsync(out->cols[up.col].offsets, 0, 4, ctx->stream));
    *out_partitioned = out.release();
    return BGPU_OK;
  }

  KeySet ks{};
  std::vector<DevAllocPtr> temps;
  BGPU_TRY(make_keyset(ctx, in, key_cols, n_keys, &ks, &temps));
  const Modulo mod = make_modulo(n);

  // grid: one persistent CTA per resident slot, contiguous chunks of whole tiles
  const size_t smem = scatter_smem_bytes(n, stage_width);
  if (smem > ctx->smem_optin) return ctx->fail(BGPU_EUNSUPPORTED, "n_partitions too large for shared memory");
  BGPU_CUDA(ctx, cudaFuncSetAttribute(scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  BGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scatter_kernel, kThreads, smem));
  if (occ < 1) occ = 1;
  int64_t g = (int64_t)ctx->sm_count * occ;
  const int64_t tiles = ceil_div64(n_rows, kTile);
  if (g > tiles) g = tiles;
  const int64_t chunk_rows = ceil_div64(tiles, g) * kTile;
  g = ceil_div64(n_rows, chunk_rows);

  DevAllocPtr hist, base, offs;
  BGPU_TRY(dev_alloc(ctx, (size_t)n * g * 4, &hist));
  BGPU_TRY(dev_alloc(ctx, (size_t)n * g * 8, &base));
  BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 8, &offs));

  count_kernel<<<(int)g, kThreads, n * 4, ctx->stream>>>(ks, mod, n_rows, chunk_rows, (uint32_t*)hist->p,
                                                        out_hashes_device);
  scan_kernel<<<1, 1024, 0, ctx->stream>>>((const uint32_t*)hist->p, (uint64_t*)base->p, (uint64_t*)offs->p, n,
                                           (uint32_t)g, n_rows);
  scatter_kernel<<<(int)g, kThreads, smem, ctx->stream>>>(ks, mod, args, n_rows, chunk_rows, (const uint64_t*)base->p);
  ctx->launches += 3;
  BGPU_CUDA(ctx, cudaGetLastError());

  for (auto& up : utf8) {  // lengths -> offsets, then copy the bytes
    const Column& c = in->cols[up.col];
    Column& o = out->cols[up.col];
    int32_t first_last[2] = {0, 0};
    BGPU_CUDA(ctx, cudaMemcpyAsync(&first_last[0], c.offsets, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(&first_last[1], c.offsets + n_rows, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaMemsetAsync(o.offsets, 0, 4, ctx->stream));
    size_t tmp_bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)up.len->p, o.offsets + 1, n_rows, ctx->stream);
    DevAllocPtr tmp;
    BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
    BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)up.len->p, o.offsets + 1, n_rows,
                                                 ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t data_bytes = (int64_t)first_last[1] - first_last[0];
    DevAllocPtr data;
    BGPU_TRY(dev_alloc(ctx, (size_t)data_bytes, &data));
    o.owned.push_back(data); o.values = data->p; o.data_bytes = data_bytes;
    const int grid = (int)std::min<int64_t>(ceil_div64(n_rows, kThreads), ctx->sm_count * 8);
    copy_bytes_kernel<<<grid, kThreads, 0, ctx->stream>>>((const uint8_t*)c.values, (const int32_t*)up.src->p,
                                                         o.offsets, (uint8_t*)data->p, n_rows);
    ctx->launches += 2;
    BGPU_CUDA(ctx, cudaGetLastError());
  }

  uint64_t* h_offs = (uint64_t*)ctx->pinned;
  BGPU_CUDA(ctx, cudaMemcpyAsync(h_offs, offs->p, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (uint32_t p = 0; p <= n; p++) out_offsets[p] = h_offs[p];
  *out_partitioned = out.release();
  return BGPU_OK;
}

extern "C" int bgpu_hash_partition_host(bgpu_ctx* ctx, const struct ArrowArray* batch, const struct ArrowSchema* schema,
                                        const int32_t* key_cols, int32_t n_keys, uint32_t n_partitions,
                                        struct ArrowArray* out_partitioned, struct ArrowSchema* out_schema,
                                        uint64_t* out_offsets) {
  if (!ctx) return BGPU_EINVAL;
  bgpu_batch *in = nullptr, *out = nullptr;
  int rc = bgpu_batch_upload(ctx, batch, schema, &in);
  if (rc == BGPU_OK) rc = bgpu_hash_partition(ctx, in, key_cols, n_keys, n_partitions, &out, out_offsets, nullptr);
  if (rc == BGPU_OK) rc = bgpu_batch_download(ctx, out, 0, out->n_rows, out_partitioned, out_schema);
  bgpu_batch_destroy(out);
  bgpu_batch_destroy(in);
  return rc;
}

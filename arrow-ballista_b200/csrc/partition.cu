// K1-K3: create_hashes, partition ids + histogram, scan, stable scatter (take-by-partition).
//
// Replaces DataFusion's BatchPartitioner::partition hash arm as driven by ShuffleWriterExec
// (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-391):
//   hash = create_hashes(keys, seeds 0,0,0,0); partition = hash % n; per partition, rows in input order.
//
// Layout: the batch is cut into G contiguous row chunks, one per persistent CTA (G = resident CTAs).
//   pass 1 (count):   each CTA hashes its chunk's keys, keeps the partition ids (2 B/row) and histograms them in shared
//                     memory -> hist[p][chunk]
//   scan:             exclusive scan of hist in (partition, chunk) order -> base[p][chunk], offsets[p]
//   pass 2 (scatter): each CTA walks its chunk in tiles; a tile is ranked stably by partition (pass 1's ids, warp match +
//                     per-warp counters), every column is staged through shared memory in partition order and written
//                     out in runs (coalesced), validity bits with atomicOr.
// HBM traffic per row: key columns once for the hash, 2 + 2 B for the partition id, every column once in and once out:
// 92 B/row for the Int64-key + 4x8 B payload microbench vs 80 B/row algorithmic.
#include <cub/device/device_scan.cuh>

#include "hash.cuh"
#include "internal.h"

namespace bgpu {

#ifndef BGPU_PART_THREADS
#define BGPU_PART_THREADS 256
#endif
constexpr int kThreads = BGPU_PART_THREADS;
constexpr uint32_t kMaxPartitions = 2048;

// ---- K1: create_hashes -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) hash_kernel(const __grid_constant__ KeySet ks, int64_t n_rows,
                                                        uint64_t* __restrict__ out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x)
    out[r] = row_hash(ks, r);
}

// ---- K2: partition ids + per-chunk histogram ----------------------------------------------------------------
constexpr uint32_t kNoPartition = 0xffffffffu;
// The lanes of the warp whose id equals this lane's (ids < 2^bits; lanes with valid = false match nobody). MATCH.ANY takes
// time in proportion to the number of DISTINCT values among the lanes; one ballot per id bit costs the same whatever the
// ids are.
__device__ __forceinline__ uint32_t match_bits(uint32_t id, bool valid, int bits) {
  if (bits <= 4) {  // at most 16 distinct ids: the hardware match wins (n = 8: 1.84 ms against 2.04 ms per 1e8 rows)
    const uint32_t peers = __match_any_sync(0xffffffffu, valid ? id : kNoPartition);
    return valid ? peers : 0u;
  }
  uint32_t peers = __ballot_sync(0xffffffffu, valid);
  for (int b = 0; b < bits; b++) {
    const bool one = (id >> b) & 1u;
    const uint32_t m = __ballot_sync(0xffffffffu, one);
    peers &= one ? m : ~m;
  }
  return valid ? peers : 0u;
}
constexpr int kCountThreads = 1024;  // the chunk layout is the scatter's; the histogram pass just wants residency
__global__ void __launch_bounds__(kCountThreads)
    count_kernel(const __grid_constant__ KeySet ks, const __grid_constant__ Modulo mod, int64_t n_rows,
                 int64_t chunk_rows, uint32_t* __restrict__ hist /* [n][gridDim.x] */,
                 uint64_t* __restrict__ out_hashes, uint16_t* __restrict__ out_pids /* n <= 2048: the scatter's input */) {
  extern __shared__ uint32_t s_hist[];
  const uint32_t n = mod.n;
  for (uint32_t p = threadIdx.x; p < n; p += kCountThreads) s_hist[p] = 0;
  __syncthreads();
  const int64_t begin = (int64_t)blockIdx.x * chunk_rows;
  int64_t end = begin + chunk_rows;
  if (end > n_rows) end = n_rows;
  constexpr int U = 8;  // rows per thread in flight
  int64_t r = begin + threadIdx.x;
  if (ks.fast64) {  // single 8-byte key column: all U loads go out before the first hash
    const uint64_t* keys = (const uint64_t*)ks.cols[0].values;
    for (; r + (int64_t)(U - 1) * kCountThreads < end; r += (int64_t)U * kCountThreads) {
      uint64_t k[U];
#pragma unroll
      for (int u = 0; u < U; u++) k[u] = __ldg(keys + r + (int64_t)u * kCountThreads);
#pragma unroll
      for (int u = 0; u < U; u++) {
        const uint64_t h = hash_u64(k[u]);
        if (out_hashes) out_hashes[r + (int64_t)u * kCountThreads] = h;
        const uint32_t p = partition_of(h, mod);
        out_pids[r + (int64_t)u * kCountThreads] = (uint16_t)p;
        atomicAdd(&s_hist[p], 1u);  // (electing one lane per distinct id with a warp match was measured slower: 0.30 vs 0.20 ms)
      }
    }
  }
  for (; r < end; r += kCountThreads) {
    const uint64_t h = row_hash(ks, r);
    if (out_hashes) out_hashes[r] = h;
    const uint32_t p = partition_of(h, mod);
    out_pids[r] = (uint16_t)p;
    atomicAdd(&s_hist[p], 1u);
  }
  __syncthreads();
  for (uint32_t p = threadIdx.x; p < n; p += kCountThreads) hist[(size_t)p * gridDim.x + blockIdx.x] = s_hist[p];
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
// ONE kernel for the local repartition (bgpu_hash_partition) and for the fused repartition + exchange over peer memory
// (bgpu_partition_exchange, PEER = true: the destination of a run is the owner rank's window). Per tile of
// THREADS x ITEMS rows: the partition ids pass 1 left behind (2 B/row; no key is read or hashed again) are ranked stably
// with warp match + per-warp 16-bit counters; fixed-width columns of 4, 8 and 16 bytes are staged ALL AT ONCE with cp.async
// (LDGSTS: global -> shared at the row's sorted position, no registers held), so a tile exposes one DRAM latency for all
// of them; then every column is written out in partition runs, the destination row of a sorted position computed once.
// The remaining kinds (SIDE = true: 1/2-byte values, utf8 (length, start) pairs, bitmaps with atomics) go one at a time
// through a small side buffer while the async copies are in flight. The lean instantiation (no SIDE work) fits 64
// registers: four 256-thread CTAs per SM, so one CTA's loads overlap another's ranking and stores.
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
  int32_t stage_off;        // byte offset of this column's async stage area for a tile of ONE row (x tile rows); -1 = side buffer
};
struct AsyncCol { const void* in; void* out; int32_t width; int32_t stage_off; };
struct ScatterArgs {
  ScatterCol cols[kMaxCols];
  AsyncCol async_cols[kMaxCols];  // the cp.async-staged columns, widest first
  int8_t side_cols[kMaxCols];     // the columns with side work (bitmaps, utf8, side-buffer values)
  int32_t n_cols, n_async, n_side;
  int32_t async_width;  // sum of the widths of the async-staged columns
  int32_t side_width;   // bytes per element of the side buffer (0 = none needed)
};
constexpr uint32_t kNoPid = 0xffffffffu;

template <int BYTES>
__device__ __forceinline__ void cp_async(uint32_t smem_dst, const void* gsrc) {
  if (BYTES == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
  else if (BYTES == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_dst), "l"(gsrc) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}

// rows w*32*ITEMS + it*32 + lane of the tile: `in` points at the column's element of this thread's first row
template <int ITEMS, typename T>
__device__ __forceinline__ void stage_async(const T* __restrict__ in, uint32_t stage /* shared address */, const uint32_t (&pid)[ITEMS],
                                            const uint32_t (&pos)[ITEMS]) {
#pragma unroll
  for (int it = 0; it < ITEMS; it++)
    if (pid[it] != kNoPid) cp_async<sizeof(T)>(stage + pos[it] * (uint32_t)sizeof(T), in + it * 32);
}

// write one staged column out: sorted position j = threadIdx.x + k * THREADS goes to output row d[k] (-1 = none)
template <int THREADS, int ITEMS, typename T>
__device__ __forceinline__ void write_runs(T* __restrict__ out, const T* stage, const int64_t (&d)[ITEMS]) {
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (d[k] >= 0) out[d[k]] = stage[threadIdx.x + k * THREADS];
}
// the same into the owner ranks' windows: outs[own[k]] is the column's region in the window of the run's owner
template <int THREADS, int ITEMS, typename T>
__device__ __forceinline__ void write_runs_peer(void* const* outs, const T* stage, const int64_t (&d)[ITEMS], const int (&own)[ITEMS]) {
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (d[k] >= 0) ((T*)outs[own[k]])[d[k]] = stage[threadIdx.x + k * THREADS];
}
// the same with the output row recomputed per element (side-buffer columns: keeps registers for the loads)
template <int THREADS, int ITEMS, typename T>
__device__ __forceinline__ void write_runs(T* __restrict__ out, const T* stage, const uint16_t* pid_sorted,
                                           const int64_t* gdelta, int tile_n) {
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    const int j = threadIdx.x + k * THREADS;
    if (j < tile_n) out[(int64_t)j + gdelta[pid_sorted[j]]] = stage[j];
  }
}

// one column through the side buffer: load -> stage in partition order -> sync -> write -> sync
template <int THREADS, int ITEMS, typename T, typename Load>
__device__ __forceinline__ void scatter_side(Load load, T* __restrict__ out, T* stage, int64_t tile_base, int tile_n,
                                             const uint32_t (&pos)[ITEMS], const uint16_t* pid_sorted,
                                             const int64_t* gdelta) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kBatch = sizeof(T) >= 16 ? 4 : (sizeof(T) >= 8 ? 8 : 16);  // loads in flight per thread
  constexpr int B = kBatch < ITEMS ? kBatch : ITEMS;
#pragma unroll
  for (int b0 = 0; b0 < ITEMS; b0 += B) {
    T v[B];
#pragma unroll
    for (int k = 0; k < B; k++) {
      const int local = w * (32 * ITEMS) + (b0 + k) * 32 + lane;
      if (local < tile_n) v[k] = load(tile_base + local);
    }
#pragma unroll
    for (int k = 0; k < B; k++) {
      const int local = w * (32 * ITEMS) + (b0 + k) * 32 + lane;
      if (local < tile_n) stage[pos[b0 + k]] = v[k];
    }
  }
  __syncthreads();
  write_runs<THREADS, ITEMS, T>(out, stage, pid_sorted, gdelta, tile_n);
  __syncthreads();
}

// Scatter a bitmap (validity or boolean values): set bit d of the zero-initialised output for every set input bit.
template <int THREADS, int ITEMS>
__device__ __forceinline__ void scatter_bits(const uint8_t* __restrict__ in, uint32_t* __restrict__ out,
                                             const int64_t* gdelta, const uint16_t* pid_sorted, int64_t tile_base,
                                             int tile_n, const uint32_t (&pos)[ITEMS]) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int it = 0; it < ITEMS; it++) {
    const int local = w * (32 * ITEMS) + it * 32 + lane;
    if (local < tile_n && bit_is_set(in, tile_base + local)) {
      const int64_t d = (int64_t)pos[it] + gdelta[pid_sorted[pos[it]]];
      atomicOr(&out[d >> 5], 1u << (d & 31));
    }
  }
}

// shared memory of one scatter CTA (tile = threads * items rows); peer_ptrs = n_async * world for the fused exchange
static size_t scatter_smem_bytes(uint32_t n, int threads, int items, int async_width, int side_width, int peer_ptrs = 0) {
  const size_t tile = (size_t)threads * items;
  const size_t ns = ((size_t)n + 1) & ~(size_t)1;
  size_t b = tile * (size_t)(async_width + side_width);
  b = (b + 15) & ~(size_t)15;
  b += (size_t)n * 8 * 2 + (size_t)n * 4 + (size_t)(threads / 32) * ns * 2 + tile * 2;
  b = (b + 7) & ~(size_t)7;
  b += (size_t)peer_ptrs * 8 + (peer_ptrs ? (size_t)n : 0);
  return b + 16;
}

template <int THREADS, bool SIDE> struct ScatterCtas {  // resident CTAs per SM the register budget is set for
  static constexpr int value = THREADS >= 1024 ? 1 : (THREADS >= 512 ? (SIDE ? 1 : 2) : (SIDE ? 2 : 4));
};

template <int THREADS, int ITEMS, bool SIDE, bool PEER>
__global__ void __launch_bounds__(THREADS, ScatterCtas<THREADS, SIDE>::value)
    scatter_kernel(const __grid_constant__ ScatterArgs args, const uint16_t* __restrict__ pids /* pass 1's partition ids */,
                   uint32_t n, int64_t n_rows, int64_t chunk_rows, const uint64_t* __restrict__ base /* [n][gridDim.x] */,
                   void* const* __restrict__ peer_out /* PEER: [n_async][world] column regions in the owners' windows */, int world) {
  constexpr int kTile = THREADS * ITEMS;
  constexpr int W = THREADS / 32;
  extern __shared__ __align__(16) unsigned char smem[];
  // shared layout: async stage areas (column c at stage_off * kTile, 16 B aligned) | side buffer | gdelta i64[n] |
  //   chunk_base u64[n] | tile_off u32[n] | warp_cnt u16[W][ns] | pid_sorted u16[kTile] | PEER: out ptrs [n_async][world], owner u8[n]
  const uint32_t ns = (n + 1) & ~1u;
  unsigned char* sp = smem;
  unsigned char* stage = sp;              sp += (size_t)kTile * args.async_width;
  void* side = sp;                        if (SIDE) sp += (size_t)kTile * args.side_width;
  sp = (unsigned char*)(((uintptr_t)sp + 15) & ~(uintptr_t)15);
  int64_t* gdelta = (int64_t*)sp;         sp += (size_t)n * 8;
  uint64_t* chunk_base = (uint64_t*)sp;   sp += (size_t)n * 8;
  uint32_t* tile_off = (uint32_t*)sp;     sp += (size_t)n * 4;
  uint16_t* warp_cnt = (uint16_t*)sp;     sp += (size_t)W * ns * 2;
  uint16_t* pid_sorted = (uint16_t*)sp;   sp += (size_t)kTile * 2;
  sp = (unsigned char*)(((uintptr_t)sp + 7) & ~(uintptr_t)7);
  void** s_out = (void**)sp;              if (PEER) sp += (size_t)args.n_async * world * 8;
  uint8_t* s_owner = (uint8_t*)sp;
  __shared__ uint32_t s_scan[W];
  (void)side; (void)s_out; (void)s_owner;

  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1;
  const int64_t begin = (int64_t)blockIdx.x * chunk_rows;
  int64_t end = begin + chunk_rows;
  if (end > n_rows) end = n_rows;
  for (uint32_t p = threadIdx.x; p < n; p += THREADS) chunk_base[p] = base[(size_t)p * gridDim.x + blockIdx.x];
  for (uint32_t i = threadIdx.x; i < W * ns / 2; i += THREADS) ((uint32_t*)warp_cnt)[i] = 0;
  if (PEER) {
    for (int i = threadIdx.x; i < args.n_async * world; i += THREADS) s_out[i] = peer_out[i];
    for (uint32_t p = threadIdx.x; p < n; p += THREADS) s_owner[p] = (uint8_t)(p % (uint32_t)world);
  }
  uint16_t* wc = warp_cnt + (size_t)w * ns;  // this warp's counters
  const int id_bits = 32 - __clz(n > 1 ? n - 1 : 1);

  // the next tile's partition ids are loaded while this tile is moved: the ranking never waits on DRAM
  uint32_t next_pid[ITEMS];
  auto prefetch_pids = [&](int64_t tb) {
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
      const int64_t r = tb + w * (32 * ITEMS) + it * 32 + lane;
      next_pid[it] = r < end ? (uint32_t)__ldg(pids + r) : kNoPid;
    }
  };
  prefetch_pids(begin);
  __syncthreads();

  for (int64_t tile_base = begin; tile_base < end; tile_base += kTile) {
    const int tile_n = (int)((end - tile_base) < kTile ? (end - tile_base) : kTile);
    // 1. stable rank within this warp's segment
    uint32_t pid[ITEMS], pos[ITEMS];
#pragma unroll
    for (int it = 0; it < ITEMS; it++) pid[it] = next_pid[it];
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
      const uint32_t peers = match_bits(pid[it], pid[it] != kNoPid, id_bits);
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if (lane == leader) {  // (lanes without a row have no peers: leader = -1)
        old = wc[pid[it]];
        wc[pid[it]] = (uint16_t)(old + __popc(peers));
      }
      old = __shfl_sync(0xffffffffu, old, leader);
      pos[it] = old + __popc(peers & lt_mask);
      __syncwarp();
    }
    __syncthreads();
    // 2. per partition: exclusive scan over warps, tile totals, then exclusive scan over partitions
    for (uint32_t p = threadIdx.x; p < n; p += THREADS) {
      uint32_t run = 0;
#pragma unroll
      for (int ww = 0; ww < W; ww++) {
        const uint32_t c = warp_cnt[(size_t)ww * ns + p];
        warp_cnt[(size_t)ww * ns + p] = (uint16_t)run;
        run += c;
      }
      tile_off[p] = run;  // the count; the scan below turns it into the offset
    }
    __syncthreads();
    {  // block exclusive scan of the counts in tile_off[0..n), in place; each thread owns a contiguous slice
      const uint32_t per = (n + THREADS - 1) / THREADS;
      const uint32_t p0 = threadIdx.x * per;
      uint32_t sum = 0;
      for (uint32_t k = 0; k < per; k++)
        if (p0 + k < n) sum += tile_off[p0 + k];
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
          const uint32_t cnt = tile_off[p];
          tile_off[p] = run;
          gdelta[p] = (int64_t)chunk_base[p] - (int64_t)run;  // sorted position -> output row (PEER: row in the owner's window)
          chunk_base[p] += cnt;
          run += cnt;
        }
    }
    __syncthreads();
    // 3. position of every row in the partition-ordered tile; all async columns go in flight
#pragma unroll
    for (int it = 0; it < ITEMS; it++) {
      if (pid[it] != kNoPid) {
        pos[it] += tile_off[pid[it]] + wc[pid[it]];
        pid_sorted[pos[it]] = (uint16_t)pid[it];
      }
    }
    {
      const int64_t r0 = tile_base + w * (32 * ITEMS) + lane;  // this thread's rows: r0 + it * 32
      const uint32_t stage_s = (uint32_t)__cvta_generic_to_shared(stage);
      for (int c = 0; c < args.n_async; c++) {
        const AsyncCol& ac = args.async_cols[c];
        const uint32_t st = stage_s + (uint32_t)ac.stage_off * kTile;
        if (ac.width == 8) stage_async<ITEMS, uint64_t>((const uint64_t*)ac.in + r0, st, pid, pos);
        else if (ac.width == 16) stage_async<ITEMS, uint4>((const uint4*)ac.in + r0, st, pid, pos);
        else stage_async<ITEMS, uint32_t>((const uint32_t*)ac.in + r0, st, pid, pos);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    prefetch_pids(tile_base + kTile);
    __syncwarp();  // this warp's counters were read above: clear them for the next tile (no other warp touches them before
    for (uint32_t i = lane; i < ns / 2; i += 32) ((uint32_t*)wc)[i] = 0;  // the next tile's barrier)
    // 4. bitmaps and side-buffer columns while the async copies land
    if (SIDE) {
      __syncthreads();
      for (int sc = 0; sc < args.n_side; sc++) {
        const ScatterCol& col = args.cols[args.side_cols[sc]];
        if (col.in_valid) scatter_bits<THREADS, ITEMS>(col.in_valid, col.out_valid, gdelta, pid_sorted, tile_base, tile_n, pos);
        if (col.kind == K_BOOL) {
          scatter_bits<THREADS, ITEMS>((const uint8_t*)col.in, (uint32_t*)col.out, gdelta, pid_sorted, tile_base, tile_n, pos);
        } else if (col.kind == K_UTF8 || col.kind == K_BINARY) {
          const int32_t* offs = col.in_offsets;
          scatter_side<THREADS, ITEMS, int32_t>([=](int64_t r) { return offs[r + 1] - offs[r]; }, col.out_len, (int32_t*)side,
                                       tile_base, tile_n, pos, pid_sorted, gdelta);
          scatter_side<THREADS, ITEMS, int32_t>([=](int64_t r) { return offs[r]; }, col.out_src, (int32_t*)side, tile_base, tile_n,
                                       pos, pid_sorted, gdelta);
        } else if (col.stage_off < 0) {
          switch (col.width) {
            case 16: {
              const uint4* in = (const uint4*)col.in;
              scatter_side<THREADS, ITEMS, uint4>([=](int64_t r) { return in[r]; }, (uint4*)col.out, (uint4*)side, tile_base,
                                         tile_n, pos, pid_sorted, gdelta);
            } break;
            case 8: {
              const uint64_t* in = (const uint64_t*)col.in;
              scatter_side<THREADS, ITEMS, uint64_t>([=](int64_t r) { return in[r]; }, (uint64_t*)col.out, (uint64_t*)side,
                                            tile_base, tile_n, pos, pid_sorted, gdelta);
            } break;
            case 4: {
              const uint32_t* in = (const uint32_t*)col.in;
              scatter_side<THREADS, ITEMS, uint32_t>([=](int64_t r) { return in[r]; }, (uint32_t*)col.out, (uint32_t*)side,
                                            tile_base, tile_n, pos, pid_sorted, gdelta);
            } break;
            case 2: {
              const uint16_t* in = (const uint16_t*)col.in;
              scatter_side<THREADS, ITEMS, uint16_t>([=](int64_t r) { return in[r]; }, (uint16_t*)col.out, (uint16_t*)side,
                                            tile_base, tile_n, pos, pid_sorted, gdelta);
            } break;
            default: {
              const uint8_t* in = (const uint8_t*)col.in;
              scatter_side<THREADS, ITEMS, uint8_t>([=](int64_t r) { return in[r]; }, (uint8_t*)col.out, (uint8_t*)side, tile_base,
                                           tile_n, pos, pid_sorted, gdelta);
            } break;
          }
        }
      }
    }
    // 5. the async-staged columns: sorted position j = threadIdx.x + k * THREADS -> output row d[k], once for all columns
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    {
      int64_t d[ITEMS];
      int own[ITEMS];
#pragma unroll
      for (int k = 0; k < ITEMS; k++) {
        const int j = threadIdx.x + k * THREADS;
        d[k] = -1;
        if (PEER) own[k] = 0;
        if (j < tile_n) {
          const uint32_t p = pid_sorted[j];
          d[k] = (int64_t)j + gdelta[p];
          if (PEER) own[k] = s_owner[p];
        }
      }
      for (int c = 0; c < args.n_async; c++) {
        const AsyncCol& ac = args.async_cols[c];
        const unsigned char* st = stage + (size_t)ac.stage_off * kTile;
        if (PEER) {  // a run of one partition is contiguous in the owner's window: a warp's stores fill whole NVLink packets
          void* const* outs = s_out + c * world;
          if (ac.width == 8) write_runs_peer<THREADS, ITEMS, uint64_t>(outs, (const uint64_t*)st, d, own);
          else if (ac.width == 16) write_runs_peer<THREADS, ITEMS, uint4>(outs, (const uint4*)st, d, own);
          else write_runs_peer<THREADS, ITEMS, uint32_t>(outs, (const uint32_t*)st, d, own);
        } else {
          if (ac.width == 8) write_runs<THREADS, ITEMS, uint64_t>((uint64_t*)ac.out, (const uint64_t*)st, d);
          else if (ac.width == 16) write_runs<THREADS, ITEMS, uint4>((uint4*)ac.out, (const uint4*)st, d);
          else write_runs<THREADS, ITEMS, uint32_t>((uint32_t*)ac.out, (const uint32_t*)st, d);
        }
      }
    }
    // (the next tile's barriers order these reads before its writes of gdelta / pid_sorted / the stage areas)
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
                                  uint64_t* out_hashes_device) try {
  if (!ctx || !batch || !key_cols || !out_hashes_device) return BGPU_EINVAL;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  KeySet ks{};
  std::vector<DevAllocPtr> temps;
  BGPU_TRY(make_keyset(ctx, batch, key_cols, n_keys, &ks, &temps));
  if (batch->n_rows == 0) return BGPU_OK;
  const int grid = (int)std::min<int64_t>(ceil_div64(batch->n_rows, kThreads), ctx->sm_count * 8);
  hash_kernel<<<grid, kThreads, 0, ctx->stream>>>(ks, batch->n_rows, out_hashes_device);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

// scatter geometries (threads per CTA, rows per thread): tile = threads x items rows
struct ScatterGeom { int threads, items; };
constexpr ScatterGeom kGeomNarrow{256, 2};   //  512-row tiles: rows too wide for two 1024-row CTAs per SM
constexpr ScatterGeom kGeomSmall{256, 4};    // 1024-row tiles, up to four CTAs per SM: few partitions (runs stay long)
constexpr ScatterGeom kGeomMedium{512, 4};   // 2048-row tiles, two CTAs per SM: many partitions
constexpr ScatterGeom kGeomLarge{1024, 4};   // 4096-row tiles, one CTA per SM
constexpr size_t kTwoCtasBudget = 112 * 1024;  // dynamic shared memory that still lets two CTAs share an SM

struct ScatterLaunch {
  const ScatterArgs* args;
  const uint16_t* pids;
  uint32_t n;
  int64_t n_rows, chunk_rows;
  const uint64_t* base;
  void* const* peer_out;  // fused exchange only
  int world;
};
template <int THREADS, int ITEMS, bool SIDE, bool PEER>
static int scatter_run(bgpu_ctx* ctx, bool launch, int64_t g, size_t smem, const ScatterLaunch& l, int* occ) {
  BGPU_CUDA(ctx, raise_max_dynamic_smem((const void*)scatter_kernel<THREADS, ITEMS, SIDE, PEER>, ctx->device, smem));
  if (!launch) {
    BGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, scatter_kernel<THREADS, ITEMS, SIDE, PEER>, THREADS, smem));
    return BGPU_OK;
  }
  scatter_kernel<THREADS, ITEMS, SIDE, PEER><<<(int)g, THREADS, smem, ctx->stream>>>(*l.args, l.pids, l.n, l.n_rows, l.chunk_rows, l.base,
                                                                                   l.peer_out, l.world);
  return BGPU_OK;
}
// launch = false: *occ = resident CTAs per SM of the instantiation (the grid and the chunks follow from it)
static int scatter_dispatch(bgpu_ctx* ctx, ScatterGeom geom, bool side, bool peer, bool launch, int64_t g, size_t smem,
                            const ScatterLaunch& l, int* occ) {
#define BGPU_SCATTER_CASE(T, I)                                                                      \
  if (geom.threads == T && geom.items == I) {                                                        \
    if (peer) return scatter_run<T, I, false, true>(ctx, launch, g, smem, l, occ);                   \
    if (side) return scatter_run<T, I, true, false>(ctx, launch, g, smem, l, occ);                   \
    return scatter_run<T, I, false, false>(ctx, launch, g, smem, l, occ);                            \
  }
  BGPU_SCATTER_CASE(256, 2)
  BGPU_SCATTER_CASE(256, 4)
  BGPU_SCATTER_CASE(512, 4)
  BGPU_SCATTER_CASE(1024, 4)
#undef BGPU_SCATTER_CASE
  return ctx->fail(BGPU_EINVAL, "internal: unknown scatter geometry");
}
static bool parse_geometry(const char* e, ScatterGeom* geom) {
  int t = 0, it = 0;
  if (!e || sscanf(e, "%d,%d", &t, &it) != 2) return false;
  if (!((t == 256 && (it == 2 || it == 4)) || (t == 512 && it == 4) || (t == 1024 && it == 4))) return false;
  *geom = ScatterGeom{t, it};
  return true;
}

// The repartition proper: pass 1 is the caller's (`pass1` launches a kernel that leaves one 16-bit partition id per row and
// hist[p][chunk], the chunk layout being the scatter's), then scan + scatter of every column. out_offsets == nullptr: the
// caller does not want the partition offsets on the host (no read-back, no synchronisation unless a utf8 column of unknown
// extent needs one).
int bgpu::partition_rows(bgpu_ctx* ctx, const bgpu_batch* in, uint32_t n_partitions, const PartitionPass1& pass1,
                         bgpu_batch** out_partitioned, uint64_t* out_offsets) {
  *out_partitioned = nullptr;
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
      if (c.kind == K_DICT32) o.owned.insert(o.owned.end(), c.owned.begin(), c.owned.end());  // shares the dictionary
    }
  }
  // Stage plan: fixed 4/8/16-byte columns are staged together with cp.async (widest first, so every area is aligned to
  // its element size); everything else shares a side buffer as wide as its widest element. Tile = 256 x items rows:
  // the largest of 16/8/4 items per thread that leaves room for two CTAs per SM (one loads while the other writes);
  // if the schema is too wide for that, columns are demoted to the side buffer (widest first) at 16 items.
  std::vector<size_t> async_cols;
  int side_width = 0;
  for (size_t i = 0; i < in->cols.size(); i++) {
    const Column& c = in->cols[i];
    args.cols[i].stage_off = -1;
    if (c.kind == K_UTF8 || c.kind == K_BINARY) side_width = std::max(side_width, 4);
    else if (c.kind == K_BOOL) continue;
    else if (c.width >= 4 && ((uintptr_t)c.values % (size_t)c.width) == 0) async_cols.push_back(i);
    else side_width = std::max(side_width, (int)c.width);
  }
  std::stable_sort(async_cols.begin(), async_cols.end(), [&](size_t a, size_t b) { return in->cols[a].width > in->cols[b].width; });
  // Geometry: what decides the speed is the length of a run (tile rows / n): below ~32 rows (256 B of an 8-byte column) the
  // head and tail sectors of every run are partial writes and the time climbs steeply (1e8 rows, ms: 128-row runs 1.85,
  // 64: 1.96, 32: 2.2, 16: 2.8, 8: 4.5, 4: 7.4); small tiles in turn keep four CTAs per SM busy. n <= 16: 1024-row tiles,
  // n <= 64: 2048-row tiles x two CTAs, more: 4096-row tiles; rows too wide for that take the next smaller tile, and a
  // schema too wide even for 512-row tiles demotes columns to the side buffer, widest first.
  auto need_for = [&](ScatterGeom gm) {
    int aw = 0;
    for (size_t i : async_cols) aw += in->cols[i].width;
    return scatter_smem_bytes(n, gm.threads, gm.items, aw, side_width);
  };
  ScatterGeom geom = kGeomSmall;
  bool forced = parse_geometry(getenv("BGPU_PART_GEOMETRY"), &geom);  // tuning knob: "threads,items"
  for (;;) {
    if (!forced) {
      geom = kGeomNarrow;
      if (need_for(kGeomSmall) <= kTwoCtasBudget) geom = kGeomSmall;
      if (n > 16 && need_for(kGeomMedium) <= kTwoCtasBudget) geom = kGeomMedium;
      if (n > 64 && need_for(kGeomLarge) <= ctx->smem_optin) geom = kGeomLarge;
    }
    const size_t need = need_for(geom);
    const size_t budget = geom.threads >= 1024 ? ctx->smem_optin : std::min<size_t>(kTwoCtasBudget, ctx->smem_optin);
    if (need <= budget) break;
    if (async_cols.empty()) {
      if (need <= ctx->smem_optin) break;
      if (geom.threads == 256 && geom.items == 2) return ctx->fail(BGPU_EUNSUPPORTED, "n_partitions too large for shared memory");
      geom = kGeomNarrow;  // the per-partition arrays alone overflow the larger tile
      forced = true;
      continue;
    }
    side_width = std::max(side_width, (int)in->cols[async_cols.front()].width);  // demote the widest async column
    async_cols.erase(async_cols.begin());
  }
  {
    int off = 0;
    args.n_async = 0;
    for (size_t i : async_cols) {
      args.cols[i].stage_off = off;
      args.async_cols[args.n_async++] = AsyncCol{args.cols[i].in, args.cols[i].out, in->cols[i].width, off};
      off += in->cols[i].width;
    }
    args.async_width = off;
    args.side_width = side_width;
    args.n_side = 0;
    for (size_t i = 0; i < in->cols.size(); i++) {
      const Column& c = in->cols[i];
      if (c.validity || c.kind == K_BOOL || c.kind == K_UTF8 || c.kind == K_BINARY || args.cols[i].stage_off < 0)
        args.side_cols[args.n_side++] = (int8_t)i;
    }
  }

  if (n_rows == 0) {
    if (out_offsets)
      for (uint32_t p = 0; p <= n; p++) out_offsets[p] = 0;
    for (auto& up : utf8) BGPU_CUDA(ctx, cudaMemsetAsync(out->cols[up.col].offsets, 0, 4, ctx->stream));
    *out_partitioned = out.release();
    return BGPU_OK;
  }

  // grid: one persistent CTA per resident slot, contiguous chunks of whole tiles
  const size_t smem = scatter_smem_bytes(n, geom.threads, geom.items, args.async_width, args.side_width);
  const bool side = args.n_side > 0;
  const int64_t kTile = (int64_t)geom.threads * geom.items;
  ScatterLaunch l{&args, nullptr, n, n_rows, 0, nullptr, nullptr, 1};
  int occ = 1;
  BGPU_TRY(scatter_dispatch(ctx, geom, side, false, false, 0, smem, l, &occ));
  if (occ < 1) occ = 1;
  int64_t g = (int64_t)ctx->sm_count * occ;
  const int64_t tiles = ceil_div64(n_rows, kTile);
  if (g > tiles) g = tiles;
  const int64_t chunk_rows = ceil_div64(tiles, g) * kTile;
  g = ceil_div64(n_rows, chunk_rows);

  DevAllocPtr hist, base, offs, pids;
  BGPU_TRY(dev_alloc(ctx, (size_t)n * g * 4, &hist));
  BGPU_TRY(dev_alloc(ctx, (size_t)n * g * 8, &base));
  BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 8, &offs));
  BGPU_TRY(dev_alloc(ctx, (size_t)n_rows * 2, &pids));

  BGPU_TRY(pass1(g, chunk_rows, (uint32_t*)hist->p, (uint16_t*)pids->p));
  scan_kernel<<<1, 1024, 0, ctx->stream>>>((const uint32_t*)hist->p, (uint64_t*)base->p, (uint64_t*)offs->p, n,
                                           (uint32_t)g, n_rows);
  l.pids = (const uint16_t*)pids->p; l.chunk_rows = chunk_rows; l.base = (const uint64_t*)base->p;
  BGPU_TRY(scatter_dispatch(ctx, geom, side, false, true, g, smem, l, nullptr));
  ctx->launches += 3;
  BGPU_CUDA(ctx, cudaGetLastError());

  for (auto& up : utf8) {  // lengths -> offsets, then copy the bytes
    const Column& c = in->cols[up.col];
    Column& o = out->cols[up.col];
    // the partitioned column holds exactly the bytes the input rows reference: when the input knows an upper bound of
    // that (data_bytes) the output is sized by it and nothing is read back; borrowed device buffers of unknown extent
    // cost one round trip for their first and last offset
    int32_t first_last[2] = {0, 0};
    const bool bounded = c.data_bytes >= 0;
    if (!bounded) {
      BGPU_CUDA(ctx, cudaMemcpyAsync(&first_last[0], c.offsets, 4, cudaMemcpyDeviceToHost, ctx->stream));
      BGPU_CUDA(ctx, cudaMemcpyAsync(&first_last[1], c.offsets + n_rows, 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    BGPU_CUDA(ctx, cudaMemsetAsync(o.offsets, 0, 4, ctx->stream));
    size_t tmp_bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)up.len->p, o.offsets + 1, n_rows, ctx->stream);
    DevAllocPtr tmp;
    BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
    BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)up.len->p, o.offsets + 1, n_rows,
                                                 ctx->stream));
    if (!bounded) BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t data_bytes = bounded ? c.data_bytes : (int64_t)first_last[1] - first_last[0];
    DevAllocPtr data;
    BGPU_TRY(dev_alloc(ctx, (size_t)data_bytes, &data));
    o.owned.push_back(data); o.values = data->p; o.data_bytes = data_bytes;
    const int grid = (int)std::min<int64_t>(ceil_div64(n_rows, kThreads), ctx->sm_count * 8);
    copy_bytes_kernel<<<grid, kThreads, 0, ctx->stream>>>((const uint8_t*)c.values, (const int32_t*)up.src->p,
                                                         o.offsets, (uint8_t*)data->p, n_rows);
    ctx->launches += 2;
    BGPU_CUDA(ctx, cudaGetLastError());
  }

  if (out_offsets) {
    uint64_t* h_offs = (uint64_t*)ctx->pinned;
    BGPU_CUDA(ctx, cudaMemcpyAsync(h_offs, offs->p, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (uint32_t p = 0; p <= n; p++) out_offsets[p] = h_offs[p];
  }
  *out_partitioned = out.release();
  return BGPU_OK;
}

extern "C" int bgpu_hash_partition(bgpu_ctx* ctx, const bgpu_batch* in, const int32_t* key_cols, int32_t n_keys,
                                   uint32_t n_partitions, bgpu_batch** out_partitioned, uint64_t* out_offsets,
                                   uint64_t* out_hashes_device) try {
  if (!ctx || !in || !key_cols || !out_partitioned || !out_offsets) return BGPU_EINVAL;
  *out_partitioned = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (n_partitions == 0) return ctx->fail(BGPU_EINVAL, "n_partitions must be > 0");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  KeySet ks{};
  std::vector<DevAllocPtr> temps;
  if (in->n_rows) BGPU_TRY(make_keyset(ctx, in, key_cols, n_keys, &ks, &temps));
  const Modulo mod = make_modulo(n_partitions);
  const int64_t n_rows = in->n_rows;
  return partition_rows(ctx, in, n_partitions,
                        [&](int64_t g, int64_t chunk_rows, uint32_t* hist, uint16_t* pids) {
                          count_kernel<<<(int)g, kCountThreads, n_partitions * 4, ctx->stream>>>(ks, mod, n_rows, chunk_rows, hist,
                                                                                               out_hashes_device, pids);
                          return BGPU_OK;
                        },
                        out_partitioned, out_offsets);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" int bgpu_hash_partition_host(bgpu_ctx* ctx, const struct ArrowArray* batch, const struct ArrowSchema* schema,
                                        const int32_t* key_cols, int32_t n_keys, uint32_t n_partitions,
                                        struct ArrowArray* out_partitioned, struct ArrowSchema* out_schema,
                                        uint64_t* out_offsets) try {
  if (!ctx) return BGPU_EINVAL;
  bgpu_batch *in = nullptr, *out = nullptr;
  int rc = bgpu_batch_upload(ctx, batch, schema, &in);
  if (rc == BGPU_OK) rc = bgpu_hash_partition(ctx, in, key_cols, n_keys, n_partitions, &out, out_offsets, nullptr);
  if (rc == BGPU_OK) rc = bgpu_batch_download(ctx, out, 0, out->n_rows, out_partitioned, out_schema);
  bgpu_batch_destroy(out);
  bgpu_batch_destroy(in);
  return rc;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

#include "peer_exchange.cuh"

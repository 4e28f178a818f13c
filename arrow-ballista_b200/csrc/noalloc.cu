// bgpu_partition_offsets_noalloc: the partition offsets of BatchPartitioner's hash arm (hash % n histogram + exclusive scan,
// the same numbers bgpu_hash_partition returns in out_offsets) for callers that may not allocate: code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc. No counterpart in the
// reference.
//
// The per-chunk histograms bgpu_hash_partition keeps in a pool allocation live here in a file-scope __device__ array of a
// fixed, modest size (kScratchWords * 4 bytes = 2 MiB + 256 B on every GPU that loads the library, until exit). The entry
// point calls no cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine: one cudaMemsetAsync of the
// ticket word and one kernel launch on the ctx stream. Every call on a device shares the array: calls that overlap on the
// GPU (two streams, two contexts of one device) race - serialise them.
#include <map>
#include <mutex>

#include "hash.cuh"
#include "internal.h"

namespace bgpu {
constexpr uint32_t kNoAllocMaxPartitions = 512;  // stated maximum of n_partitions
constexpr uint32_t kNoAllocMaxChunks = 1024;     // row chunks (= CTAs) per call
constexpr uint32_t kNoAllocHeaderWords = 64;     // word 0 = ticket of finished CTAs (must start at zero); 256-byte header
constexpr size_t kScratchWords = (size_t)kNoAllocHeaderWords + (size_t)kNoAllocMaxPartitions * kNoAllocMaxChunks;
constexpr int kNoAllocThreads = 1024;

// In HOST code this name is only a stand-in for the symbol, not a device address: the kernel gets the address that
// cudaGetSymbolAddress returns (noalloc_scratch below), taken once per device with that device current.
static __device__ __align__(256) uint32_t g_noalloc_scratch[kScratchWords];

// One launch: every CTA histograms its chunk of keys (create_hashes of a single non-null Int64 column, seeds 0,0,0,0,
// hash % n) in shared memory and stores it as hist[p][chunk]; the CTA that draws the last ticket adds the chunks up
// per partition and writes the exclusive scan: out_offsets[p] = first row of partition p, out_offsets[n] = n_rows.
__global__ void __launch_bounds__(kNoAllocThreads)
    partition_offsets_noalloc_kernel(const uint64_t* __restrict__ keys, int64_t n_rows, int64_t chunk_rows,
                                     const __grid_constant__ Modulo mod, uint32_t* __restrict__ scratch,
                                     uint64_t* __restrict__ out_offsets) {
  __shared__ uint32_t s_hist[kNoAllocMaxPartitions];
  __shared__ uint64_t s_total[kNoAllocMaxPartitions];
  __shared__ bool s_last;
  const uint32_t n = mod.n, g = gridDim.x;
  uint32_t* hist = scratch + kNoAllocHeaderWords;
  for (uint32_t p = threadIdx.x; p < n; p += kNoAllocThreads) s_hist[p] = 0;
  __syncthreads();
  const int64_t begin = (int64_t)blockIdx.x * chunk_rows;
  const int64_t end = begin + chunk_rows < n_rows ? begin + chunk_rows : n_rows;
  constexpr int U = 4;  // keys per thread in flight
  int64_t r = begin + threadIdx.x;
  for (; r + (int64_t)(U - 1) * kNoAllocThreads < end; r += (int64_t)U * kNoAllocThreads) {
    uint64_t k[U];
#pragma unroll
    for (int u = 0; u < U; u++) k[u] = __ldg(keys + r + (int64_t)u * kNoAllocThreads);
#pragma unroll
    for (int u = 0; u < U; u++) atomicAdd(&s_hist[partition_of(hash_u64(k[u]), mod)], 1u);
  }
  for (; r < end; r += kNoAllocThreads) atomicAdd(&s_hist[partition_of(hash_u64(__ldg(keys + r)), mod)], 1u);
  __syncthreads();
  for (uint32_t p = threadIdx.x; p < n; p += kNoAllocThreads) hist[(size_t)p * g + blockIdx.x] = s_hist[p];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&scratch[0], 1u) == g - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last CTA: one warp per partition sums that partition's chunks
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t p = (uint32_t)warp; p < n; p += kNoAllocThreads / 32) {
    uint64_t sum = 0;
    for (uint32_t c = (uint32_t)lane; c < g; c += 32) sum += __ldcg(&hist[(size_t)p * g + c]);
#pragma unroll
    for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
    if (lane == 0) s_total[p] = sum;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t run = 0;
    for (uint32_t p = 0; p < n; p++) {
      out_offsets[p] = run;
      run += s_total[p];
    }
    out_offsets[n] = run;
    scratch[0] = 0;  // leave the ticket as the next call needs it (the entry point clears it as well)
  }
}

// hash % n constants (as bgpu_hash_partition derives them)
static Modulo make_modulo_noalloc(uint32_t n) {
  Modulo m{};
  m.n = n;
  m.pow2 = (n & (n - 1)) == 0;
  m.mask = n - 1;
  m.small = n < 65536;
  m.magic = n > 1 ? (UINT64_MAX / n + 1) : 0;
  m.two32_mod = (uint32_t)((1ULL << 32) % n);
  return m;
}

// The scratch array's device address on `device` (current when this is called): looked up once per device.
static int noalloc_scratch(bgpu_ctx* ctx, uint32_t** out) {
  static std::mutex mu;
  static std::map<int, uint32_t*> by_device;
  std::lock_guard<std::mutex> lk(mu);
  auto it = by_device.find(ctx->device);
  if (it == by_device.end()) {
    uint32_t* p = nullptr;
    BGPU_CUDA(ctx, cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (!p) return ctx->fail(BGPU_ECUDA, "cudaGetSymbolAddress returned a null scratch address");
    it = by_device.emplace(ctx->device, p).first;
  }
  *out = it->second;
  return BGPU_OK;
}
}  // namespace bgpu

using namespace bgpu;

extern "C" int bgpu_partition_offsets_noalloc(bgpu_ctx* ctx, const int64_t* keys_device, int64_t n_rows,
                                              uint32_t n_partitions, uint64_t* out_offsets_device) try {
  if (!ctx || !out_offsets_device || n_rows < 0 || (!keys_device && n_rows > 0)) return BGPU_EINVAL;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (n_partitions == 0) return ctx->fail(BGPU_EINVAL, "n_partitions must be > 0");
  if (n_partitions > kNoAllocMaxPartitions)
    return ctx->fail(BGPU_EINVAL, "bgpu_partition_offsets_noalloc: n_partitions " + std::to_string(n_partitions) +
                                      " exceeds the static scratch's maximum of " + std::to_string(kNoAllocMaxPartitions) +
                                      " (use bgpu_hash_partition)");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  uint32_t* scratch = nullptr;
  BGPU_TRY(noalloc_scratch(ctx, &scratch));
  // chunks of whole 4096-row tiles, at most kNoAllocMaxChunks of them (rows beyond that make the chunks longer)
  const int64_t tiles = std::max<int64_t>(1, ceil_div64(n_rows, 4096));
  int64_t g = std::min<int64_t>(tiles, std::min<int64_t>((int64_t)ctx->sm_count * 2, kNoAllocMaxChunks));
  const int64_t chunk_rows = ceil_div64(tiles, g) * 4096;
  g = std::max<int64_t>(1, ceil_div64(n_rows, chunk_rows));
  const Modulo mod = make_modulo_noalloc(n_partitions);
  BGPU_CUDA(ctx, cudaMemsetAsync(scratch, 0, 4, ctx->stream));  // the ticket must start at zero
  partition_offsets_noalloc_kernel<<<(int)g, kNoAllocThreads, 0, ctx->stream>>>((const uint64_t*)keys_device, n_rows, chunk_rows, mod,
                                                                              scratch, out_offsets_device);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

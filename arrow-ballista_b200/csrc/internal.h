// libbgpu internals: context, device-resident columns, error plumbing.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include <map>
#include <mutex>
#include <utility>

#include "../../include/bgpu.h"

#include <exception>
#include <new>

struct bgpu_ctx;
namespace bgpu {
// What a C ABI entry point returns when its body threw (function-try-blocks around every compute entry point): an
// allocation that failed or a container that refused a size is an error code with a message, never an abort.
int abi_guard(bgpu_ctx* ctx, const std::exception& e);
}  // namespace bgpu

// cudaFuncAttributeMaxDynamicSharedMemorySize of a statically compiled kernel, only ever raised: the attribute belongs to
// the kernel (per device), not to a launch, and tasks on different threads launch the same kernel with different
// amounts - lowering it for one between another's set and launch would fail that launch. It is a cap, not a reservation.
inline cudaError_t raise_max_dynamic_smem(const void* kernel, int device, size_t smem) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, size_t> current;
  std::lock_guard<std::mutex> lk(mu);
  auto key = std::make_pair(kernel, device);
  auto it = current.find(key);
  if (it != current.end() && it->second >= smem) return cudaSuccess;
  const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e == cudaSuccess) current[key] = smem;
  return e;
}

struct bgpu_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool owns_stream = true;
  bool has_device = false;  // false: creation failed (no GPU); every compute call fails loudly
  int sm_count = 148;
  size_t smem_optin = 0;
  std::string err;
  uint64_t launches = 0;
  // lifetime: the creator holds one reference (dropped by bgpu_ctx_destroy); every device allocation, batch and
  // aggregate holds one too, so destroying handles in any order (e.g. at interpreter shutdown) is safe
  std::atomic<int> refs{1};
  // small pinned scratch for device->host scalars (partition offsets, group counts)
  void* pinned = nullptr;
  size_t pinned_bytes = 0;
  // pinned staging blocks of the streamed input (ingest.cu), kept for the life of the context: pinning memory costs
  // milliseconds per block, an aggregate per stage must not pay it again
  std::mutex pinned_pool_mu;
  std::vector<std::pair<void*, size_t>> pinned_pool;
  // opt-in accounting of GPU time per class (bgpu_ctx_enable_times / bgpu_ctx_times: the h2d_time / kernel_time /
  // d2h_time metrics of an operator): every instrumented entry point brackets its stream work with an event pair
  bool times_on = false;
  int times_depth = 0;                       // only the outermost scope of a call chain records
  double times_ms[3] = {0.0, 0.0, 0.0};      // resolved so far: h2d, kernel, d2h
  struct TimedSpan { cudaEvent_t a, b; int cls; };
  std::vector<TimedSpan> times_open;         // recorded, not yet read back
  std::vector<cudaEvent_t> times_pool;       // events to reuse

  int fail(int code, const std::string& msg) {
    err = msg;
    return code;
  }
};

#define BGPU_CUDA(ctx, call)                                                                         \
  do {                                                                                               \
    cudaError_t _e = (call);                                                                         \
    if (_e != cudaSuccess)                                                                           \
      return (ctx)->fail(BGPU_ECUDA, std::string(#call) + ": " + cudaGetErrorString(_e) + " at " +  \
                                         __FILE__ + ":" + std::to_string(__LINE__));                 \
  } while (0)

#define BGPU_TRY(expr)        \
  do {                        \
    int _rc = (expr);         \
    if (_rc != BGPU_OK) return _rc; \
  } while (0)

// GPU time of the enclosing scope's stream work, accounted to a class (no-op unless bgpu_ctx_enable_times was called).
enum TimeClass : int { T_H2D = 0, T_KERNEL = 1, T_D2H = 2 };
struct GpuTimeScope {
  bgpu_ctx* ctx;
  cudaEvent_t a = nullptr;
  int cls;
  cudaStream_t stream;
  GpuTimeScope(bgpu_ctx* c, int cls_, cudaStream_t st = nullptr);
  ~GpuTimeScope();
  GpuTimeScope(const GpuTimeScope&) = delete;
  GpuTimeScope& operator=(const GpuTimeScope&) = delete;
};

// a pinned host block of at least `bytes` from the context's pool (cudaHostAlloc on a miss) / back into it
int ctx_pinned_get(bgpu_ctx* ctx, size_t bytes, void** out);
void ctx_pinned_put(bgpu_ctx* ctx, void* p, size_t bytes);
void ctx_retain(bgpu_ctx* ctx);
void ctx_release(bgpu_ctx* ctx);  // tears the context down when the last reference goes

// A device allocation freed (stream-ordered) when the last column referencing it goes away.
struct DevAlloc {
  void* p = nullptr;
  size_t bytes = 0;
  bgpu_ctx* ctx = nullptr;
  bool owned = true;
  std::shared_ptr<void> keep;  // something else to keep alive with the columns that hold this (a peer-window lease)
  ~DevAlloc() {
    if (p && owned && ctx) {
      cudaSetDevice(ctx->device);
      cudaFreeAsync(p, ctx->stream);
    }
    if (ctx) ctx_release(ctx);
  }
};
using DevAllocPtr = std::shared_ptr<DevAlloc>;

// Allocate `bytes` (at least 16, 256-byte aligned by the pool) on the ctx stream.
int dev_alloc(bgpu_ctx* ctx, size_t bytes, DevAllocPtr* out);

enum ColKind : int32_t { K_FIXED = 0, K_BOOL = 1, K_UTF8 = 2, K_BINARY = 3, K_DICT32 = 4 };

struct Column {
  std::string name;
  std::string format;  // Arrow C format string of this column (index type for dictionaries)
  int64_t flags = ARROW_FLAG_NULLABLE;
  ColKind kind = K_FIXED;
  int32_t width = 0;           // K_FIXED byte width; K_DICT32: 4
  void* values = nullptr;      // fixed values / bool bits / utf8 data / dictionary keys
  int32_t* offsets = nullptr;  // utf8/binary: length+1 entries, offsets[0] may be non-zero
  uint8_t* validity = nullptr; // nullptr = all valid
  int64_t length = 0;
  int64_t null_count = 0;      // -1 = unknown
  int64_t data_bytes = 0;      // utf8/binary: upper bound of the offsets, i.e. bytes addressable behind `values`; -1 = unknown
  int64_t data_first = 0;      // utf8/binary: first byte behind `values` that is really allocated (uploads of sliced arrays
                               // keep their offsets and point `values` before the copied range)
  std::shared_ptr<Column> dict;  // K_DICT32: dictionary values column
  int64_t dict_id = -1;          // K_DICT32 read from / written to an IPC stream: the dictionary's id in that stream
  std::vector<DevAllocPtr> owned;  // keeps buffers alive (shared with other batches when passed through)
};

struct bgpu_batch {
  bgpu_ctx* ctx = nullptr;
  int64_t n_rows = 0;
  std::vector<Column> cols;
  explicit bgpu_batch(bgpu_ctx* c) : ctx(c) { ctx_retain(c); }
  ~bgpu_batch() {
    cols.clear();
    ctx_release(ctx);
  }
  bgpu_batch(const bgpu_batch&) = delete;
  bgpu_batch& operator=(const bgpu_batch&) = delete;
};

// Parse one Arrow C schema node into kind/width/format. Returns BGPU_EUNSUPPORTED for types off the path.
int parse_field(bgpu_ctx* ctx, const struct ArrowSchema* s, Column* out);

// Plain-old-data view of a column handed to kernels.
struct KCol {
  const void* values;
  const int32_t* offsets;
  const uint8_t* validity;
  const uint64_t* dict_hashes;  // K_DICT32 key columns: create_hashes of the dictionary values
  int32_t kind;
  int32_t width;
};

// dictionary.cu: dictionary columns on the move (decode to values, move, re-encode with identity indices)
namespace bgpu {
// partition.cu: scan + stable scatter of every column of `in` by the 16-bit partition ids `pass1` leaves behind
// (pass1(grid, chunk_rows, hist[p][chunk], ids[row]) launches the kernel that computes them on ctx->stream)
using PartitionPass1 = std::function<int(int64_t, int64_t, uint32_t*, uint16_t*)>;
int partition_rows(bgpu_ctx* ctx, const bgpu_batch* in, uint32_t n_partitions, const PartitionPass1& pass1,
                   bgpu_batch** out_partitioned, uint64_t* out_offsets);
bool batch_has_dictionary(const bgpu_batch* b);
int dict_decode_column(bgpu_ctx* ctx, const Column& c, int64_t n, Column* out);
int dict_decode_batch(bgpu_ctx* ctx, const bgpu_batch* in, std::unique_ptr<bgpu_batch>* out, std::vector<size_t>* dict_cols);
int dict_identity_encode(bgpu_ctx* ctx, bgpu_batch* b, size_t i, const Column& proto, const uint64_t* d_offsets, uint32_t n_parts);
}  // namespace bgpu

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

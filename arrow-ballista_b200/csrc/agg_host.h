// Host-side state of one grouped aggregate (bgpu_agg), shared by aggregate.cu (plan building) and
// aggregate_run.cu (update / emit).
#pragma once
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "agg_plan.h"

namespace bgpu { struct JitModule; struct Interner; }

struct PhysAcc {
  int32_t op = 0;
  int32_t value = -1;  // index into bgpu_agg::values
  int32_t words = 1;
  int32_t out_off = 0;  // first word inside a group's accumulator record
};
struct LogicalAgg {
  int32_t fn = 0;
  int32_t phys[2] = {-1, -1};  // Sum: [sum]; Count: [count]; Avg: [count, sum]
  std::string name;
  std::string sum_format, result_format;  // Arrow formats of the state sum and of the final value
  bool decimal = false;
  int scale_up = 0;  // Avg(decimal): result scale - sum scale
};

constexpr int kFusedRungs = 4;  // private 512x2; shared-atomic 3 CTAs/SM, 1 CTA/SM, 1 CTA/SM without key cache

struct bgpu_agg {
  bgpu_ctx* ctx = nullptr;
  int32_t mode = 0;
  bool final_mode = false;
  std::vector<Column> fields;       // raw input schema
  std::vector<int32_t> group_cols;  // indices into `fields`
  std::vector<int32_t> key_off, key_width;  // packed-key layout per group column
  int32_t key_words = 1;
  // interned layout (agg_key_layout): group columns that enter the packed key as 4-byte ids of an Interner
  bool interned_mode = false;
  bool key_too_long = false;  // the last pass met a string key beyond its inline field (check_common_flags)
  std::vector<uint8_t> key_interned;
  std::vector<std::shared_ptr<bgpu::Interner>> interners;
  std::vector<bgpu::HValue> values;  // distinct aggregate arguments; a value that extends another follows it
  std::vector<bool> value_from_prev;
  std::vector<bgpu::HFilterTerm> filter;
  std::vector<PhysAcc> phys;  // sorted by value
  std::vector<LogicalAgg> aggs;
  // group table
  uint32_t max_groups = 0, n_groups = 0;
  // slots: the table (key inline); acc: one record of acc_stride u64 per group
  DevAllocPtr slots, acc, counters /* u32 [0]=n_groups, [1..4]=flags */, seen, scratch;
  uint32_t acc_stride = 1;
  bool seen_tracking = false;
  bool async_ready = false;  // the current fused rung has committed a batch: updates may be launched without reading back
  int fused_rung = 0;  // first rung of the fused ladder still worth trying; kFusedRungs = two-pass only
  uint32_t private_gcap = 4;  // groups per thread the private rung starts with (doubles when a batch needs more)
  uint32_t shared_gcap = 128; // group capacity the multi-CTA shared rung starts with: few groups = one copy per warp
  // dictionary-typed group columns are keyed by value: the aggregate keeps one canonical dictionary per column
  // (host copy; uploaded at emit) and maps every batch dictionary onto it
  struct DictState {
    std::unordered_map<std::string, int32_t> index;
    std::vector<std::string> values;
    std::string value_format = "u";
    ColKind value_kind = K_UTF8;
    std::shared_ptr<bgpu::Interner> interner;  // exact path: the canonical dictionary, built on the device (interner.cu)
    std::shared_ptr<bgpu::Interner> spare;     // the last stream's interner, emptied by bgpu_agg_reset (capacity kept)
    std::shared_ptr<Column> canon_dev;  // fast path: the first dictionary seen IS the canonical one (device resident)
    std::shared_ptr<Column> last;  // last batch dictionary seen (kept alive so its identity cannot be recycled)
    DevAllocPtr last_remap;        // nullptr = identity
  };
  std::vector<DictState> dicts;
  // what bgpu_agg_create_expr was given, byte for byte: with a batch shape, the key of the process-wide cache of
  // run-time specialised kernels (aggregate_run.cu specialise)
  std::string plan_key;
  bool emitted = false;
  // introspection for tests / bench: which path the last update took ("fused_private", "fused_shared2", "fused_shared1",
  // "shared", "global")
  std::string last_path;
  // device time of the fused kernel launches so far (CUDA events on the ctx stream): bench.py's roofline numerator
  cudaEvent_t ev[2] = {nullptr, nullptr};
  double hot_kernel_ms = 0.0;

  // Pass-through mode of a Partial aggregate (aggregate_run.cu): when the first batch's distinct-key estimate says the
  // pass would barely reduce (most rows are their own group), input rows are emitted as state rows - key, sum = the
  // argument, count = 1 - and the FinalPartitioned side does the only aggregation, after the exchange. What DataFusion
  // later added as skip_partial_aggregation; results are identical (a Partial aggregate never promised full reduction).
  uint64_t radix_batches = 0;  // batches whose rows were moved into table-bucket order first (radix pre-pass)
  bool passthrough = false;
  bool any_committed = false;  // some batch of this stream has been aggregated into the table
  std::vector<std::unique_ptr<bgpu_batch>> pass_chunks;

  // streamed host input (ingest.cu): staging ring, device buffers, launches the host has not looked at yet
  std::shared_ptr<void> ingest;
  int (*ingest_drain)(bgpu_agg*) = nullptr;  // seal + submit + poll everything pending (set by ingest.cu)
  void (*ingest_reset)(bgpu_agg*) = nullptr; // forget errors and statistics of the last stream

  explicit bgpu_agg(bgpu_ctx* c) : ctx(c) { ctx_retain(c); }
  ~bgpu_agg() {
    for (cudaEvent_t e : ev)
      if (e) cudaEventDestroy(e);
    ctx_release(ctx);
  }
  uint32_t* d_n_groups() { return (uint32_t*)counters->p; }
  int32_t* d_flags() { return (int32_t*)counters->p + 1; }
  bgpu::GroupTable table() {
    return bgpu::GroupTable{(uint64_t*)slots->p, max_groups * 2 - 1, max_groups, d_n_groups(), d_flags(),
                            (uint64_t*)acc->p, acc_stride, key_words == 1 ? 2u : 4u};
  }
};

// ---- binding of one batch to the runtime-specialised kernels ------------------------------------------------------
// One slot per distinct batch column the kernels read. Everything in JitShape is baked into the generated source;
// the pointers go into JitArgs.
struct JitSlot {
  int32_t batch_col = -1;
  int32_t kind = 0;   // ColKind
  int32_t width = 0;  // fixed width
  int32_t vk = 0;     // ValKind when read as a value / filter operand
  bool nullable = false, remap = false;
  bool key_use = false, value_use = false, filter_use = false;
};
struct JitShape {
  std::vector<JitSlot> slots;
  std::vector<int32_t> key_slot;             // per group column
  std::vector<int32_t> filter_slot;          // per filter term
  std::vector<bool> value_nullable;          // per value (in this batch)
  std::vector<int32_t> acc_alias, acc_word0; // per physical accumulator
  int32_t W = 0;                             // accumulator words per group
  bool has_seen = false;
  int32_t T = 256, R = 4, min_ctas = 1;
  // which entry points to emit: bit 0 bgpu_fused, bit 1 bgpu_intern + bgpu_acc_shared + bgpu_acc_global,
  // bit 2 bgpu_fused_shared
  int32_t kernels = 3;
  bool any_validity = false;
  bool row32 = false;                        // row indices fit 32 bits (with room for two grid strides past the end)
};
// slot of batch column `batch_col` (created on first use); -1 when the table is full
int32_t jit_slot(JitShape* sh, int32_t batch_col);
// CUDA source of the plan-specific part for (aggregate, shape): the generated `Plan` + extern "C" entry points
// bgpu_fused, bgpu_intern, bgpu_acc_shared, bgpu_acc_global. Compiled behind agg_jit_prelude().
int agg_generate_source(const bgpu_agg* g, const JitShape& sh, std::string* out);
const char* agg_jit_prelude();

int agg_launch_grid(bgpu_ctx* ctx, int64_t n, int per_block);
int agg_key_layout(bgpu_agg* g, bool interned);
// One update whose outcome is read later (streamed input): the fused launch is followed by a device-side conditional
// commit, the status words arrive in pinned memory once `done` has completed.
struct AggAsyncReq {
  uint32_t* d_status = nullptr;        // device, 8 words
  volatile uint32_t* h_status = nullptr;  // pinned host, 8 words: {committed, n_groups, counters[1..6]}
  cudaEvent_t k0 = nullptr, k1 = nullptr, done = nullptr;
  bool launched = false;               // false: the update ran synchronously (two-pass path, first batch of a shape)
};
// update_batch / merge_batch; req != nullptr asks for the asynchronous form where the current path allows it
int agg_update_impl(bgpu_agg* g, const bgpu_batch* batch, AggAsyncReq* req);
// After req->done: take the launch's outcome. *replay = true: nothing was committed, run the batch again synchronously
// (agg_update_replay).
int agg_async_result(bgpu_agg* g, AggAsyncReq* req, bool* replay);
int agg_update_replay(bgpu_agg* g, const bgpu_batch* batch);
// (Re)allocate the table and accumulators for `new_max` groups, keeping the first `keep` groups.
int agg_resize(bgpu_agg* g, uint32_t new_max, uint32_t keep);

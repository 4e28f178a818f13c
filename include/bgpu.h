/* libbgpu — C ABI of the B200-native aggregation + shuffle hot path for Ballista.
 *
 * This is the drop-in boundary: the entry points a Rust `GpuShuffleWriterExec` /
 * `GpuAggregateExec` would bind through `extern "C"` (see INTEGRATION.md for the Rust stub). Every entry
 * point names the reference interface it replaces. Paths are relative to /root/reference.
 *
 * Conventions (mirror the reference's contracts):
 *  - every call returns 0 on success, a BGPU_E* category otherwise; the message is in bgpu_last_error(ctx).
 *    Nothing aborts or throws across the ABI (reference: Result<_, DataFusionError>, shuffle_writer.rs:451-453).
 *  - inputs are *borrowed* Arrow C-Data structs valid for the call; outputs are library-allocated Arrow C-Data
 *    whose release callback frees them; opaque handles are freed by the matching *_destroy.
 *  - one bgpu_ctx per task (CUDA stream + memory pool); no global mutable state; contexts are independent and
 *    may be used from different threads concurrently (reference: DedicatedExecutor runs many tasks per
 *    process, executor_server.rs:1027-1032).
 *  - there is no CPU fallback: without a CUDA device every compute call fails with BGPU_ENODEVICE.
 */
#ifndef BGPU_H
#define BGPU_H
#include <stdint.h>

#include "bgpu_arrow_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

#define BGPU_OK 0
#define BGPU_EINVAL 1       /* bad argument / plan (DataFusionError::Plan / Internal) */
#define BGPU_EUNSUPPORTED 2 /* type or shape not handled on the GPU path (DataFusionError::NotImplemented) */
#define BGPU_ECUDA 3        /* CUDA runtime failure (DataFusionError::Execution) */
#define BGPU_ENODEVICE 4    /* no usable CUDA device */
#define BGPU_EOVERFLOW 5    /* arithmetic overflow the reference reports as an error (ArrowError::ComputeError) */
#define BGPU_ENCCL 6        /* exchange failure; maps to BallistaError::FetchFailed (shuffle_reader.rs:697-704) */

typedef struct bgpu_ctx bgpu_ctx;     /* one per task: device, stream, pool, last error */
typedef struct bgpu_batch bgpu_batch; /* a RecordBatch resident in device memory */
typedef struct bgpu_agg bgpu_agg;     /* a GroupedHashAggregateStream's state */

/* ---- context ------------------------------------------------------------------------------------------ */
int bgpu_ctx_create(int device, bgpu_ctx** out);
/* same, but all work is ordered on a caller-owned cudaStream_t (e.g. the harness's stream, so its CUDA events
 * time this library's kernels); NULL = the legacy default stream */
int bgpu_ctx_create_on_stream(int device, void* cuda_stream, bgpu_ctx** out);
void bgpu_ctx_destroy(bgpu_ctx* ctx);
const char* bgpu_last_error(const bgpu_ctx* ctx); /* never NULL; "" when no error */
int bgpu_ctx_synchronize(bgpu_ctx* ctx);
void* bgpu_ctx_stream(bgpu_ctx* ctx); /* the cudaStream_t all work of this ctx is ordered on */
const char* bgpu_version(void);
/* number of kernels this ctx has launched so far (bench.py's gpu_launches) */
uint64_t bgpu_ctx_kernel_launches(const bgpu_ctx* ctx);
/* Operator time metrics (what a GPU node adds next to the reference's write_time / repart_time,
 * /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:139-163): with accounting enabled every entry point
 * of this ctx brackets its stream work with a CUDA event pair, and bgpu_ctx_times returns the GPU milliseconds so far as
 * out3_ms = {h2d_time (bgpu_batch_upload, the streamed input's chunk copies, the IPC reader's upload + decode),
 * kernel_time (aggregate, repartition, exchange, concat, sort, IPC framing), d2h_time (bgpu_batch_download, the IPC
 * writer's copy-out)}; reset != 0 zeroes the sums. bgpu_ctx_times waits for the work it accounts for; call it when no
 * other call on the ctx is running. Off by default (two event records per call). */
int bgpu_ctx_enable_times(bgpu_ctx* ctx, int32_t on);
int bgpu_ctx_times(bgpu_ctx* ctx, double* out3_ms, int32_t reset);

/* ---- batches: RecordBatch <-> device -------------------------------------------------------------------
 * The struct-typed ArrowArray/ArrowSchema pair is a RecordBatch exactly as arrow-rs exports one
 * (format "+s", one child per column). */
/* host -> device copy (one cudaMemcpyAsync per buffer on the ctx stream) */
int bgpu_batch_upload(bgpu_ctx* ctx, const struct ArrowArray* batch, const struct ArrowSchema* schema,
                      bgpu_batch** out);
/* borrow buffers that already live in device memory (array->buffers[] are device pointers); array offsets
 * must be 0. The caller keeps them alive until bgpu_batch_destroy. */
int bgpu_batch_wrap_device(bgpu_ctx* ctx, const struct ArrowArray* batch, const struct ArrowSchema* schema,
                           bgpu_batch** out);
/* device -> host copy into library-owned host memory; rows [row_offset, row_offset+num_rows) */
int bgpu_batch_download(bgpu_ctx* ctx, const bgpu_batch* batch, int64_t row_offset, int64_t num_rows,
                        struct ArrowArray* out, struct ArrowSchema* out_schema);
/* export device pointers (no copy); valid until the batch is destroyed */
int bgpu_batch_export_device(bgpu_ctx* ctx, const bgpu_batch* batch, struct ArrowArray* out,
                             struct ArrowSchema* out_schema);
int64_t bgpu_batch_num_rows(const bgpu_batch* batch);
int32_t bgpu_batch_num_columns(const bgpu_batch* batch);
/* bytes of the batch's Arrow buffers (validity, offsets, values; utf8 data by its known bound): what write_stream_to_disk
 * adds up as num_bytes (`batch.get_array_memory_size()`, /root/reference/ballista/core/src/utils.rs:203-207) */
int64_t bgpu_batch_memory_size(const bgpu_batch* batch);
void bgpu_batch_destroy(bgpu_batch* batch);

/* ---- create_hashes ---------------------------------------------------------------------------------------
 * Replaces datafusion_common::hash_utils::create_hashes(arrays, &RandomState::with_seeds(0,0,0,0), buf)
 * as called by BatchPartitioner (call site shuffle_writer.rs:336-346; restated SURVEY.md A.1).
 * out_hashes_device: device pointer to n_rows u64. */
int bgpu_create_hashes(bgpu_ctx* ctx, const bgpu_batch* batch, const int32_t* key_cols, int32_t n_keys,
                       uint64_t* out_hashes_device);

/* ---- BatchPartitioner::partition, hash arm -----------------------------------------------------------------
 * Replaces `partitioner.partition(batch, |partition, out_batch| ...)` (shuffle_writer.rs:346-391):
 * partition = create_hashes(keys) % n_partitions; every column is gathered per partition in input order
 * (arrow `take`). The result is ONE batch whose rows are grouped by partition ascending and stable within a
 * partition; partition p is rows [out_offsets[p], out_offsets[p+1]) (empty partitions have equal offsets —
 * the caller skips them as the reference does, shuffle_writer.rs:349).
 * out_offsets: HOST pointer to n_partitions+1 u64. out_hashes_device: optional device pointer (n_rows u64). */
int bgpu_hash_partition(bgpu_ctx* ctx, const bgpu_batch* batch, const int32_t* key_cols, int32_t n_keys,
                        uint32_t n_partitions, bgpu_batch** out_partitioned, uint64_t* out_offsets,
                        uint64_t* out_hashes_device);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The partition offsets of the hash arm
 * above - out_offsets_device[p] = first row of partition p in the partitioned order, out_offsets_device[n_partitions] =
 * n_rows; exactly bgpu_hash_partition's out_offsets - for ONE non-null 8-byte key column (Int64 / UInt64 / Timestamp /
 * Float64 bits) given as a plain device pointer. keys_device (n_rows x 8 bytes) and out_offsets_device
 * ((n_partitions + 1) x 8 bytes) are the caller's device memory. The call allocates nothing (no cudaMalloc / cudaFree /
 * cudaMallocAsync, no allocating library routine): its per-chunk histograms live in a file-scope __device__ array of
 * 2 MiB + 256 B that every GPU which loads the library holds until exit. LIMIT: n_partitions <= 512 (more: BGPU_EINVAL
 * with a message). Every call on a device shares that array: calls that overlap on the GPU race - serialise them. Work is
 * ordered on the ctx stream; the result is ready after bgpu_ctx_synchronize (or later work on that stream). */
int bgpu_partition_offsets_noalloc(bgpu_ctx* ctx, const int64_t* keys_device, int64_t n_rows, uint32_t n_partitions,
                                   uint64_t* out_offsets_device);
/* Host-buffer convenience used by the operator layer: upload + partition + download in one call. */
int bgpu_hash_partition_host(bgpu_ctx* ctx, const struct ArrowArray* batch, const struct ArrowSchema* schema,
                             const int32_t* key_cols, int32_t n_keys, uint32_t n_partitions,
                             struct ArrowArray* out_partitioned, struct ArrowSchema* out_schema,
                             uint64_t* out_offsets);

/* ---- AggregateExec / GroupedHashAggregateStream ------------------------------------------------------------
 * Replaces AggregateExec(mode=Partial|FinalPartitioned) (plan shape planner.rs:376-427, proto
 * datafusion.proto:1405-1411,1437-1450; restated SURVEY.md A.3). */
#define BGPU_AGG_MODE_PARTIAL 0           /* AggregateMode::PARTIAL */
#define BGPU_AGG_MODE_FINAL 1             /* AggregateMode::FINAL */
#define BGPU_AGG_MODE_FINAL_PARTITIONED 2 /* AggregateMode::FINAL_PARTITIONED */
#define BGPU_AGG_MODE_SINGLE 3            /* AggregateMode::SINGLE (raw rows -> final values) */
#define BGPU_AGG_MIN 0 /* AggregateFunction numbering, datafusion.proto:631-636 */
#define BGPU_AGG_MAX 1
#define BGPU_AGG_SUM 2
#define BGPU_AGG_AVG 3
#define BGPU_AGG_COUNT 4

/* input_schema: schema of the *raw* input of the Partial aggregate (AggregateExecNode.input_schema,
 * datafusion.proto:1444-1445) — also required in Final modes, where the batches fed to update() are state
 * batches (group columns first, then state columns in aggregate order: Sum -> sum; Count -> count Int64;
 * Avg -> count UInt64, sum). group_cols / agg_cols index input_schema; agg_cols[i] = -1 means COUNT(*).
 * Integer arguments of AVG are cast to Float64 (the planner's coercion). agg_names (may be NULL) name the
 * output columns: Partial emits "<name>[sum]" / "<name>[count]" state columns, Final modes emit "<name>". */
int bgpu_agg_create(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                    const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                    const int32_t* agg_cols, const char* const* agg_names, int32_t n_agg, bgpu_agg** out);
/* Same, with the aggregate arguments given as physical expressions and an optional fused row filter.
 * DataFusion evaluates each aggregate's input expressions inside the aggregate stream (the plan prints
 * `aggr=[SUM(lineitem.l_extendedprice * Int64(1))]`, planner.rs:379), and Q1's `l_shipdate <= date` FilterExec
 * sits directly below the Partial aggregate (benchmarks/queries/q1.sql); both are evaluated per row inside the
 * aggregate kernels instead of materialising intermediate columns.
 * Expressions are postfix programs over the raw input schema (Partial/Single modes only):
 *   COLUMN a            push input column a
 *   LITERAL t,p,s,lo,hi push a literal of type t (BGPU_T_*); Decimal128(p,s) / Int64 / Date32 in lo,hi (i128,
 *                       two's complement, low word first); Float64 as its bit pattern in lo
 *   ADD SUB MUL         arrow-arith numeric.rs semantics: Decimal128 operands are rescaled to max(s1,s2) for +/-
 *                       (result precision max(s1,s2)+max(p1-s1,p2-s2)+1), multiply adds scales (precision p1+p2+1),
 *                       both capped at 38, and every step is i128-checked: overflow is BGPU_EOVERFLOW (the reference
 *                       returns ArrowError::ComputeError). Int64 literals next to decimals are coerced to
 *                       Decimal128(20,0) as DataFusion's type coercion does. Int64 arithmetic wraps; Float64 is IEEE.
 *   LT LE GT GE EQ NE   comparisons (filter only);  AND  conjunction (filter only)
 * A null operand makes the value null; SUM/AVG skip null values, COUNT counts non-null ones; a null or false
 * filter drops the row. agg_expr_offsets[i] .. agg_expr_offsets[i+1] delimit aggregate i's program inside
 * `nodes` (an empty program = COUNT(*)). */
#define BGPU_X_COLUMN 0
#define BGPU_X_LITERAL 1
#define BGPU_X_ADD 2
#define BGPU_X_SUB 3
#define BGPU_X_MUL 4
#define BGPU_X_LT 5
#define BGPU_X_LE 6
#define BGPU_X_GT 7
#define BGPU_X_GE 8
#define BGPU_X_EQ 9
#define BGPU_X_NE 10
#define BGPU_X_AND 11
#define BGPU_T_INT64 0
#define BGPU_T_FLOAT64 1
#define BGPU_T_DECIMAL128 2
#define BGPU_T_DATE32 3
typedef struct bgpu_expr_node {
  int32_t op;        /* BGPU_X_* */
  int32_t a;         /* COLUMN: column index; LITERAL: BGPU_T_* */
  int32_t precision; /* LITERAL Decimal128 */
  int32_t scale;
  uint64_t lo, hi;   /* LITERAL value */
} bgpu_expr_node;
int bgpu_agg_create_expr(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                         const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                         const bgpu_expr_node* nodes, const int32_t* agg_expr_offsets,
                         const char* const* agg_names, int32_t n_agg, const bgpu_expr_node* filter,
                         int32_t n_filter, bgpu_agg** out);
/* update_batch (Partial/Single) or merge_batch (Final*) with one device-resident batch */
int bgpu_agg_update(bgpu_agg* agg, const bgpu_batch* batch);
/* The operator's input side as the reference drives it: one call per HOST RecordBatch of the child stream, 8192 rows by
 * default (/root/reference/ballista/core/src/config.rs:180-182; the stream loop of
 * /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:342-346), from as many threads at once as the task
 * has input partitions (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:183-216). Partial / Single
 * modes. The call is THREAD-SAFE and returns once the batch's buffers have been read (borrowed for the call): the
 * calling thread packs the columns the aggregate reads into a pinned staging chunk in a narrow wire form
 * (frame-of-reference integers of 0/1/2/4/8 bytes - Decimal128 values that fit 64 bits included -, utf8 as narrow
 * lengths + bytes); full chunks are uploaded on a copy stream, widened back to Arrow buffers on the device and
 * aggregated by one kernel launch per chunk without a host round trip; results are bit-identical to bgpu_batch_upload +
 * bgpu_agg_update per batch. `schema` may be NULL (the batch is then trusted to match input_schema). bgpu_agg_emit and
 * bgpu_agg_update first complete everything pushed so far (callers join their pushing threads before); errors found
 * late (a launch that reports overflow) surface from the next push / flush / emit. Dictionary columns:
 * BGPU_EUNSUPPORTED (take bgpu_batch_upload + bgpu_agg_update).
 * bgpu_agg_flush: submit the partly filled chunk and wait for every launch's outcome.
 * bgpu_agg_ingest_stats (introspection): out8 = {rows, batches, wire bytes copied host->device, Arrow bytes they stand
 * for, chunks, launches not read back, launches replayed, 0}. */
int bgpu_agg_push_host(bgpu_agg* agg, const struct ArrowArray* batch, const struct ArrowSchema* schema);
int bgpu_agg_flush(bgpu_agg* agg);
int bgpu_agg_ingest_stats(const bgpu_agg* agg, uint64_t* out8);
/* test hook (host only): the wire form of one fixed-width column segment as bgpu_agg_push_host packs it (which = 0: the
 * AVX2 build of the packer, 1: the plain build, 2: the AVX-512 build); out_packed has room for n * 16 bytes */
int bgpu_debug_pack_fixed_host(const char* format, const void* values, int64_t n, int32_t which, int64_t* out_base,
                               uint32_t* out_code, uint8_t* out_packed);
/* test hook (host only): the packer's one-pass route - pack against a GIVEN base and code (1, 2 or 4 bytes) and verify on the
 * way that every value fits. 1 = fits (out_packed valid), 0 = some value does not fit, negative = -error code */
int bgpu_debug_pack_checked_host(const char* format, const void* values, int64_t n, int32_t which, int64_t base, uint32_t code,
                                 uint8_t* out_packed);
/* end of input: state() (Partial) or evaluate() (Final modes and Single). Group order is unspecified, as in the
 * reference (GroupValues uses a randomly seeded table). The aggregate may not be updated afterwards. */
int bgpu_agg_emit(bgpu_agg* agg, bgpu_batch** out);
/* A new input stream over the same plan: the reference creates its AggregateExec once per task and `execute`s it per
 * partition (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:234-257). Clears the groups and the
 * accumulators; the plan, its compiled kernels, the kernel path and capacities the last stream settled on and the staging
 * buffers of bgpu_agg_push_host stay. Allowed before or after bgpu_agg_emit. */
int bgpu_agg_reset(bgpu_agg* agg);
int64_t bgpu_agg_num_groups(bgpu_agg* agg); /* groups seen so far */
/* which kernel path the last update took ("fused_private", "fused_shared2", "fused_shared1", "shared", "global", "intern");
 * introspection for tests and the bench, not part of the reference's interface */
const char* bgpu_agg_last_path(const bgpu_agg* agg);
/* device time (ms, CUDA events on the ctx stream) spent so far in this aggregate's single-pass fused kernel launches:
 * the per-launch duration bench.py's roofline line is computed from */
double bgpu_agg_hot_kernel_ms(const bgpu_agg* agg);
void bgpu_agg_destroy(bgpu_agg* agg);
/* Run-time compiled kernels are cached twice: per process (modules) and on disk (cubins, under $BGPU_CACHE_DIR, default
 * ~/.cache/bgpu; "off" disables; keyed by source + sm_100a + NVRTC version, verified byte for byte on load), so that only
 * the first process that ever sees a plan pays NVRTC's ~2 s. out4 = {NVRTC compiles, cubins loaded from disk, in-memory
 * hits, cubins stored} of this process. */
void bgpu_jit_stats(uint64_t* out4);
/* Introspection, no device needed: generate the run-time specialised kernels of this plan (arguments as
 * bgpu_agg_create_expr) for a synthetic batch shape and compile them for sm_100a with NVRTC. flags bit 0: every column
 * carries validity; bit 1: dictionary keys need a remap; bit 2: per-accumulator seen masks are tracked.
 * threads / rows_per_thread / min_ctas = launch geometry baked into the kernels. When out_prefix is not NULL the
 * source and the cubin are written to <out_prefix>.cu / <out_prefix>.cubin (for cuobjdump -sass / -res-usage). */
int bgpu_agg_debug_compile(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                           const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                           const bgpu_expr_node* nodes, const int32_t* agg_expr_offsets, int32_t n_agg,
                           const bgpu_expr_node* filter, int32_t n_filter, int32_t flags, int32_t threads,
                           int32_t rows_per_thread, int32_t min_ctas, const char* out_prefix);

/* ---- intra-node exchange (replaces shuffle files + Flight do_get + ShuffleReaderExec fan-in;
 * shuffle_writer.rs:351-385, flight_service.rs:124-166, shuffle_reader.rs:539-624) --------------------------
 * nccl_comm: an ncclComm_t created by the host harness (one rank per process/GPU). Output partition p is
 * owned by rank p % world. `partitioned`/`offsets` come from bgpu_hash_partition with the same n_partitions on
 * every rank. out_received holds this rank's partitions from all ranks, grouped by (partition, source rank). */
int bgpu_all_to_all(bgpu_ctx* ctx, void* nccl_comm, int32_t rank, int32_t world, const bgpu_batch* partitioned,
                    const uint64_t* offsets, uint32_t n_partitions, bgpu_batch** out_received);
/* The exchange of a SMALL batch (an aggregate state of a few rows: TPC-H Q1 has four) as one fixed-size collective:
 * every rank packs its batch into a 32 KB slot, ONE ncclAllGather hands every rank every slot, one kernel keeps the rows
 * of the partitions this rank owns (create_hashes(keys) % n_partitions, owner = partition % world) - the rows
 * bgpu_all_to_all(bgpu_hash_partition(batch)) delivers, ordered by (source rank, source row). Two kernels, one
 * collective and one host round trip, whatever the number of ranks. COLLECTIVE: every rank calls it at the same point;
 * when any rank's batch does not fit (more than 256 rows or 32 KB) EVERY rank gets BGPU_EUNSUPPORTED (the slots carry
 * the flag) and takes bgpu_hash_partition + bgpu_all_to_all / bgpu_partition_exchange. Dictionary columns:
 * BGPU_EUNSUPPORTED before any collective (schema-level, the same on every rank). world == 1 needs no communicator. */
int bgpu_small_exchange(bgpu_ctx* ctx, void* nccl_comm, int32_t rank, int32_t world, const bgpu_batch* batch,
                        const int32_t* key_cols, int32_t n_keys, uint32_t n_partitions, bgpu_batch** out_received);
/* NCCL plumbing for harnesses that have no ncclComm_t of their own (libnccl.so.2 is dlopen'ed; the copy the process
 * already loaded is reused): rank 0 makes a 128-byte ncclUniqueId, the harness broadcasts it by its own means (the
 * Python mirror uses torch.distributed), every rank creates its communicator on the ctx's device. */
int bgpu_nccl_unique_id(bgpu_ctx* ctx, void* out_id128);
int bgpu_nccl_comm_create(bgpu_ctx* ctx, const void* id128, int32_t rank, int32_t world, void** out_comm);
void bgpu_nccl_comm_destroy(void* comm);
/* Host-side layout of the received batch (no device needed; also what the gloo CPU tests check): all_offsets =
 * [world][n_partitions + 1] partition offsets of every rank's partitioned batch; recv_off[(k * world + s)] = first
 * row, in this rank's received batch, of its k-th owned partition (p = rank + k * world) coming from source s, with
 * one trailing total. Returns the total. */
uint64_t bgpu_exchange_plan(const uint64_t* all_offsets, int32_t rank, int32_t world, uint32_t n_partitions,
                            uint64_t* recv_off, uint32_t* n_owned);

/* ---- fused hash repartition + exchange over NVLink peer memory (K3 + K6 in one kernel) ------------------------------
 * bgpu_partition_exchange(batch) == bgpu_all_to_all(bgpu_hash_partition(batch)): replaces, inside one NVSwitch node,
 * BatchPartitioner::partition + the shuffle file write + Flight do_get + the reader fan-in
 * (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-391,
 * /root/reference/ballista/executor/src/flight_service.rs:124-166,
 * /root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:539-624). The scatter kernel stores every run of a
 * partition straight into the owner rank's receive window (cudaIpc-mapped peer memory), so rows cross NVLink once and no
 * partitioned copy is made in local HBM; NCCL carries only the counts and a completion word.
 * All three calls are COLLECTIVE over the ranks of `nccl_comm` (bgpu_peer_exchange_destroy is not).
 * bgpu_partition_exchange accepts batches of null-free fixed-width columns of 4, 8 or 16 bytes (aggregate states),
 * n_partitions <= 128; anything else returns BGPU_EUNSUPPORTED on EVERY rank (the ranks agree through the counts
 * exchange) and the caller takes bgpu_hash_partition + bgpu_all_to_all. The received batch's columns are the receive
 * window itself (no copy-out): the batch holds the window until bgpu_batch_destroy, the next exchange uses the rank's
 * second window, and while a rank still holds the results of its last two exchanges the call returns BGPU_EUNSUPPORTED
 * on every rank. out_recv_off (optional, [n_owned * world + 1]):
 * as bgpu_exchange_plan; out_sent_bytes (optional): bytes this rank stored into other ranks' windows. */
typedef struct bgpu_peer_exchange bgpu_peer_exchange;
int bgpu_peer_exchange_create(bgpu_ctx* ctx, void* nccl_comm, int32_t rank, int32_t world, uint64_t window_bytes,
                              bgpu_peer_exchange** out);
int bgpu_partition_exchange(bgpu_peer_exchange* x, const bgpu_batch* batch, const int32_t* key_cols, int32_t n_keys,
                            uint32_t n_partitions, bgpu_batch** out_received, uint64_t* out_recv_off,
                            uint64_t* out_sent_bytes);
uint64_t bgpu_peer_exchange_window_bytes(const bgpu_peer_exchange* x);
void bgpu_peer_exchange_destroy(bgpu_peer_exchange* x);

/* ---- Arrow IPC stream framing of partitioned device batches --------------------------------------
 * Replaces arrow-ipc's StreamWriter as ShuffleWriterExec drives it, one `write(batch)` per non-empty output partition
 * (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:351-399; IPCWriter,
 * /root/reference/ballista/core/src/utils.rs:63-141). A partition file is: schema message, the RecordBatch messages
 * appended to it, then the end-of-stream marker (ff ff ff ff 00 00 00 00). Metadata version V5, buffers 8-byte aligned,
 * bodies uncompressed or LZ4_FRAME. Dictionary columns: BGPU_EUNSUPPORTED.
 *
 * bgpu_ipc_schema_message: the encapsulated Schema message of a struct-typed ArrowSchema (host only; malloc'ed: free
 *   with bgpu_ipc_free(p, 0)).
 * bgpu_ipc_record_batch_metadata: the 8-byte prefix + RecordBatch metadata of one message, for callers that lay the
 *   body out themselves (host only; nodes = (length, null_count) pairs, buffers = (offset, length) pairs; malloc'ed);
 *   compression = 0 marks the body's buffers as LZ4_FRAME-compressed (each: int64 uncompressed length, then the frame).
 * bgpu_ipc_encode_partitions: one encapsulated RecordBatch message per non-empty partition p = rows
 *   [offsets[p], offsets[p + 1]) of `batch`; bodies are laid out by one kernel on the device and arrive with one copy.
 *   *out_bytes (pinned host memory; free with bgpu_ipc_free(p, 1); NULL when every partition is empty) holds the
 *   messages back to back: message p = bytes [out_msg_off[p], out_msg_off[p + 1]) (empty range = empty partition).
 *   out_msg_off has n + 1 entries. compression = 0: every buffer is LZ4_FRAME-compressed ON THE DEVICE as the reference's
 *   writer does per buffer (shuffle_writer.rs:365-376; frames of block-independent 64 KB blocks like lz4_flex's, one warp
 *   per block; a buffer that does not shrink is stored behind a -1 length) before the (smaller) copy to the host. */
int bgpu_ipc_schema_message(bgpu_ctx* ctx, const struct ArrowSchema* schema, uint8_t** out_bytes, uint64_t* out_len);
int bgpu_ipc_record_batch_metadata(bgpu_ctx* ctx, int64_t rows, const int64_t* nodes, int32_t n_nodes, const int64_t* buffers,
                                   int32_t n_buffers, int64_t body_len, int32_t compression /* -1 none, 0 LZ4_FRAME */,
                                   uint8_t** out_bytes, uint64_t* out_len);
int bgpu_ipc_encode_partitions(bgpu_ctx* ctx, const bgpu_batch* batch, const uint64_t* offsets, uint32_t n,
                               int32_t compression /* -1 none, 0 LZ4_FRAME */, uint8_t** out_bytes, uint64_t* out_msg_off);
void bgpu_ipc_free(uint8_t* bytes, int pinned);

/* ---- Arrow IPC stream reader into device batches: the inverse of the above ----------------
 * Replaces `AsyncStreamReader::{try_new, maybe_next}` (/root/reference/ballista/core/src/async_reader/mod.rs:108-258)
 * as `fetch_partition_local_inner` drives it over one shuffle partition file
 * (/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:771-785): the Schema message first, then one
 * device batch per RecordBatch message until the end-of-stream marker, a zero length or plain EOF. Reads what the
 * reference's ShuffleWriterExec writes (LZ4_FRAME buffer compression, shuffle_writer.rs:365-376: block-independent and
 * block-linked frames, stored buffers with the -1 prefix) and what bgpu_ipc_encode_partitions writes (uncompressed).
 * LZ4 blocks are decoded on the device, one decode chain per warp. `bytes` is borrowed host memory that must stay valid
 * until bgpu_ipc_reader_close; each bgpu_ipc_reader_next returns with its copies out of it completed.
 * Errors: BGPU_EINVAL for malformed streams (messages "Unable to read IPC message as schema", "Not expecting a schema
 * when messages are read" as the reference's reader words them), BGPU_EUNSUPPORTED for dictionary batches, ZSTD,
 * big-endian streams and types off the path. */
typedef struct bgpu_ipc_reader bgpu_ipc_reader;
int bgpu_ipc_reader_open(bgpu_ctx* ctx, const uint8_t* bytes, uint64_t len, bgpu_ipc_reader** out); /* host only */
int bgpu_ipc_reader_schema(bgpu_ipc_reader* reader, struct ArrowSchema* out);                       /* host only */
int bgpu_ipc_reader_next(bgpu_ipc_reader* reader, bgpu_batch** out); /* *out = NULL at the end of the stream */
void bgpu_ipc_reader_close(bgpu_ipc_reader* reader);
/* test hook (host, one lane): the LZ4 frame walker + block decoder the device kernel is built from; never called by
 * the library itself */
int bgpu_debug_lz4_frame_decode_host(bgpu_ctx* ctx, const uint8_t* frame, uint64_t len, uint8_t* out, uint64_t out_len);

/* ---- concatenation of device batches ---------------------------------------------------------------
 * Replaces `concat_batches` as CoalesceBatchesExec(target_batch_size) applies it above the shuffle reader
 * (plan pinned at /root/reference/ballista/scheduler/src/planner.rs:386-387, 409-413): rows of batches[0..n) in order,
 * all of one schema. One kernel moves every buffer (values copied, utf8 offsets rebased onto the running byte total,
 * validity / boolean bitmaps re-aligned to the running row total; a column is nullable in the result when any input
 * has nulls). Dictionary columns: BGPU_EUNSUPPORTED. */
int bgpu_batch_concat(bgpu_ctx* ctx, const bgpu_batch* const* batches, int32_t n, bgpu_batch** out);

/* ---- stable sort of a device batch: the ordered fan-in of CoalesceTasksExec --------------------------
 * When the child of a task group is ordered, the reference merges its partitions' sorted streams with `streaming_merge`
 * (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:148-173; ties go to the lower stream index). On
 * the device: bgpu_batch_concat of the partitions' batches in stream order, then this STABLE sort by the sort expressions -
 * for sorted inputs the same rows in the same order. sort_cols index the batch's columns (fixed-width incl. Decimal128 and
 * floats in IEEE total order, Boolean, Utf8 / Binary up to 4096 bytes); descending[k] / nulls_first[k] are the
 * SortOptions of key k (NULL pointers = ascending, nulls last). Dictionary-encoded KEY columns: BGPU_EUNSUPPORTED. */
int bgpu_batch_sort(bgpu_ctx* ctx, const bgpu_batch* batch, const int32_t* sort_cols, const uint8_t* descending,
                    const uint8_t* nulls_first, int32_t n_keys, bgpu_batch** out);

#ifdef __cplusplus
}
#endif
#endif

//! never compiled here (no cargo in this image). The C ABI of include/bgpu.h as Rust sees it, RAII wrappers around the
//! opaque handles, and the error mapping of SURVEY.md §8b ("Errors": Result<_, DataFusionError>, never a panic across
//! the ABI).
use std::ffi::{c_char, c_void, CStr};

use arrow::error::ArrowError;
use arrow::ffi::{FFI_ArrowArray, FFI_ArrowSchema};
use ballista_core::error::BallistaError;
use datafusion::error::{DataFusionError, Result};

#[repr(C)] pub struct BgpuCtx { _p: [u8; 0] }
#[repr(C)] pub struct BgpuBatch { _p: [u8; 0] }
#[repr(C)] pub struct BgpuAgg { _p: [u8; 0] }
#[repr(C)] pub struct BgpuIpcReader { _p: [u8; 0] }
#[repr(C)] pub struct BgpuPeerExchange { _p: [u8; 0] }

/// bgpu_expr_node: one node of a postfix expression program (include/bgpu.h BGPU_X_* / BGPU_T_*)
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct BgpuExprNode { pub op: i32, pub a: i32, pub precision: i32, pub scale: i32, pub lo: u64, pub hi: u64 }

pub const BGPU_EUNSUPPORTED: i32 = 2;
pub const BGPU_EOVERFLOW: i32 = 5;
pub const BGPU_ENCCL: i32 = 6;
pub const AGG_PARTIAL: i32 = 0; // AggregateMode numbering, datafusion.proto:1405-1411
pub const AGG_FINAL_PARTITIONED: i32 = 2;

extern "C" {
    pub fn bgpu_ctx_create(device: i32, out: *mut *mut BgpuCtx) -> i32;
    pub fn bgpu_ctx_destroy(ctx: *mut BgpuCtx);
    pub fn bgpu_last_error(ctx: *const BgpuCtx) -> *const c_char;
    pub fn bgpu_batch_upload(ctx: *mut BgpuCtx, batch: *const FFI_ArrowArray, schema: *const FFI_ArrowSchema,
                             out: *mut *mut BgpuBatch) -> i32;
    pub fn bgpu_batch_download(ctx: *mut BgpuCtx, b: *const BgpuBatch, row_offset: i64, num_rows: i64,
                               out: *mut FFI_ArrowArray, out_schema: *mut FFI_ArrowSchema) -> i32;
    pub fn bgpu_batch_num_rows(b: *const BgpuBatch) -> i64;
    pub fn bgpu_batch_destroy(b: *mut BgpuBatch);
    // BatchPartitioner::partition, hash arm (shuffle_writer.rs:336-391)
    pub fn bgpu_hash_partition(ctx: *mut BgpuCtx, batch: *const BgpuBatch, key_cols: *const i32, n_keys: i32,
                               n_partitions: u32, out: *mut *mut BgpuBatch, out_offsets: *mut u64,
                               out_hashes_device: *mut u64) -> i32;
    pub fn bgpu_hash_partition_host(ctx: *mut BgpuCtx, batch: *const FFI_ArrowArray, schema: *const FFI_ArrowSchema,
                                    key_cols: *const i32, n_keys: i32, n_partitions: u32, out: *mut FFI_ArrowArray,
                                    out_schema: *mut FFI_ArrowSchema, out_offsets: *mut u64) -> i32;
    // AggregateExec(Partial | FinalPartitioned)
    pub fn bgpu_agg_create_expr(ctx: *mut BgpuCtx, mode: i32, input_schema: *const FFI_ArrowSchema, group_cols: *const i32,
                                n_group: i32, agg_fns: *const i32, nodes: *const BgpuExprNode, agg_expr_offsets: *const i32,
                                agg_names: *const *const c_char, n_agg: i32, filter: *const BgpuExprNode, n_filter: i32,
                                out: *mut *mut BgpuAgg) -> i32;
    pub fn bgpu_agg_push_host(agg: *mut BgpuAgg, batch: *const FFI_ArrowArray, schema: *const FFI_ArrowSchema) -> i32;
    pub fn bgpu_agg_update(agg: *mut BgpuAgg, batch: *const BgpuBatch) -> i32;
    pub fn bgpu_agg_reset(agg: *mut BgpuAgg) -> i32;
    pub fn bgpu_agg_emit(agg: *mut BgpuAgg, out: *mut *mut BgpuBatch) -> i32;
    pub fn bgpu_agg_destroy(agg: *mut BgpuAgg);
    // shuffle files: StreamWriter with LZ4_FRAME (shuffle_writer.rs:365-384) and the reader (async_reader/mod.rs:108-258)
    pub fn bgpu_ipc_schema_message(ctx: *mut BgpuCtx, schema: *const FFI_ArrowSchema, out: *mut *mut u8, len: *mut u64) -> i32;
    pub fn bgpu_ipc_encode_partitions(ctx: *mut BgpuCtx, batch: *const BgpuBatch, offsets: *const u64, n: u32,
                                      compression: i32, out_bytes: *mut *mut u8, out_msg_off: *mut u64) -> i32;
    pub fn bgpu_ipc_free(bytes: *mut u8, pinned: i32);
    pub fn bgpu_ipc_reader_open(ctx: *mut BgpuCtx, bytes: *const u8, len: u64, out: *mut *mut BgpuIpcReader) -> i32;
    pub fn bgpu_ipc_reader_next(r: *mut BgpuIpcReader, out: *mut *mut BgpuBatch) -> i32;
    pub fn bgpu_ipc_reader_close(r: *mut BgpuIpcReader);
    pub fn bgpu_batch_concat(ctx: *mut BgpuCtx, batches: *const *const BgpuBatch, n: i32, out: *mut *mut BgpuBatch) -> i32;
    // ordered CoalesceTasksExec: concat in stream order + stable sort == streaming_merge of sorted children
    pub fn bgpu_batch_sort(ctx: *mut BgpuCtx, batch: *const BgpuBatch, sort_cols: *const i32, descending: *const u8,
                           nulls_first: *const u8, n_keys: i32, out: *mut *mut BgpuBatch) -> i32;
    // intra-node exchange (replaces files + Flight do_get + reader fan-in)
    pub fn bgpu_small_exchange(ctx: *mut BgpuCtx, comm: *mut c_void, rank: i32, world: i32, batch: *const BgpuBatch,
                               key_cols: *const i32, n_keys: i32, n_partitions: u32, out: *mut *mut BgpuBatch) -> i32;
    pub fn bgpu_partition_exchange(x: *mut BgpuPeerExchange, batch: *const BgpuBatch, key_cols: *const i32, n_keys: i32,
                                   n_partitions: u32, out: *mut *mut BgpuBatch, out_recv_off: *mut u64,
                                   out_sent_bytes: *mut u64) -> i32;
    pub fn bgpu_all_to_all(ctx: *mut BgpuCtx, comm: *mut c_void, rank: i32, world: i32, partitioned: *const BgpuBatch,
                           offsets: *const u64, n_partitions: u32, out: *mut *mut BgpuBatch) -> i32;
    // h2d_time / kernel_time / d2h_time of the operators that ran on a context (MetricValue::Time next to write_time)
    pub fn bgpu_ctx_enable_times(ctx: *mut BgpuCtx, on: i32) -> i32;
    pub fn bgpu_ctx_times(ctx: *mut BgpuCtx, out3_ms: *mut f64, reset: i32) -> i32;
    // callers that may not allocate (no counterpart in the reference)
    pub fn bgpu_partition_offsets_noalloc(ctx: *mut BgpuCtx, keys_device: *const i64, n_rows: i64, n_partitions: u32,
                                          out_offsets_device: *mut u64) -> i32;
}

/// One per task (SURVEY.md §8b "Threading"): a CUDA stream + pool. Send: a task may migrate between worker threads.
pub struct Ctx(pub *mut BgpuCtx);
unsafe impl Send for Ctx {}
unsafe impl Sync for Ctx {}
impl Ctx {
    pub fn new(device: i32) -> Result<Self> {
        let mut p = std::ptr::null_mut();
        let rc = unsafe { bgpu_ctx_create(device, &mut p) };
        let ctx = Ctx(p);
        ctx.check(rc)?;
        Ok(ctx)
    }
    /// rc -> DataFusionError
    pub fn check(&self, rc: i32) -> Result<()> {
        if rc == 0 { return Ok(()); }
        let msg = unsafe { CStr::from_ptr(bgpu_last_error(self.0)) }.to_string_lossy().into_owned();
        Err(match rc {
            BGPU_EUNSUPPORTED => DataFusionError::NotImplemented(msg),
            BGPU_EOVERFLOW => DataFusionError::ArrowError(ArrowError::ComputeError(msg)),
            // an exchange that failed is a lost shuffle input: the scheduler re-runs the map stage (shuffle_reader.rs:697-704)
            BGPU_ENCCL => DataFusionError::External(Box::new(BallistaError::FetchFailed(String::new(), 0, vec![], msg))),
            _ => DataFusionError::Execution(msg),
        })
    }
}
impl Drop for Ctx { fn drop(&mut self) { unsafe { bgpu_ctx_destroy(self.0) } } }

/// A RecordBatch in device memory.
pub struct DeviceBatch(pub *mut BgpuBatch);
unsafe impl Send for DeviceBatch {}
impl DeviceBatch { pub fn num_rows(&self) -> usize { unsafe { bgpu_batch_num_rows(self.0) as usize } } }
impl Drop for DeviceBatch { fn drop(&mut self) { unsafe { bgpu_batch_destroy(self.0) } } }

pub struct Agg(pub *mut BgpuAgg);
unsafe impl Send for Agg {}
unsafe impl Sync for Agg {} // bgpu_agg_push_host is thread-safe; everything else is called by one task
impl Drop for Agg { fn drop(&mut self) { unsafe { bgpu_agg_destroy(self.0) } } }

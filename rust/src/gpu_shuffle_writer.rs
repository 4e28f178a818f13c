//! never compiled here (no cargo in this image). `GpuShuffleWriterExec`: the reference's `ShuffleWriterExec`
//! (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:234-449, 459-597) with the hash arm on the device.
//! Same fields, same `ExecutionPlan` answers (schema = the input's, output partitioning = the INPUT's, :483-493), same
//! files (`<work_dir>/<job>/<stage>/<partition>/<uuid>.arrow`, IPC stream, LZ4_FRAME per buffer, :358-376), same
//! `ShuffleWritePartition` accounting (`num_bytes` = file length before `finish()`, :396-399).
//! Python mirror: arrow-ballista_b200/shuffle_writer.py.
use std::fs::File;
use std::io::Write;
use std::sync::Arc;

use arrow::ffi::FFI_ArrowSchema;
use ballista_core::serde::protobuf::ShuffleWritePartition;
use datafusion::error::{DataFusionError, Result};
use datafusion::execution::TaskContext;
use datafusion::physical_plan::expressions::Column;
use datafusion::physical_plan::{ExecutionPlan, Partitioning};

use crate::ffi::*;
use crate::gpu_aggregate::GpuAggregateExec;

#[derive(Debug)]
pub struct GpuShuffleWriterExec {
    pub job_id: String,
    pub stage_id: usize,
    pub plan: Arc<dyn ExecutionPlan>,   // a GpuAggregateExec when the state is to stay on the device
    pub work_dir: String,
    pub shuffle_output_partitioning: Option<Partitioning>,
    pub device: i32,
}

struct PartitionFile { file: File, path: String, num_batches: u64, num_rows: u64, num_bytes: u64 }

const EOS: [u8; 8] = [0xff, 0xff, 0xff, 0xff, 0, 0, 0, 0];
const LZ4_FRAME: i32 = 0;

impl GpuShuffleWriterExec {
    /// execute_shuffle_write, `Some(Partitioning::Hash(exprs, n))` arm (:328-449), for the plan shape
    /// ShuffleWriterExec(Hash) -> [CoalesceTasksExec] -> AggregateExec(Partial).
    pub async fn execute_shuffle_write(&self, input_partition: usize, task: Arc<TaskContext>) -> Result<Vec<ShuffleWritePartition>> {
        let (exprs, n) = match &self.shuffle_output_partitioning {
            Some(Partitioning::Hash(exprs, n)) => (exprs, *n),
            _ => return Err(DataFusionError::Execution("Invalid shuffle partitioning scheme".to_owned())), // :451-453
        };
        let keys: Vec<i32> = exprs.iter()
            .map(|e| e.as_any().downcast_ref::<Column>().map(|c| c.index() as i32)
                .ok_or_else(|| DataFusionError::NotImplemented("hash key that is not a column".into())))
            .collect::<Result<_>>()?;
        let agg = self.plan.as_any().downcast_ref::<GpuAggregateExec>()
            .ok_or_else(|| DataFusionError::Internal("GpuShuffleWriterExec expects a GpuAggregateExec child".into()))?;
        let ctx = Ctx::new(self.device)?;
        // the child's whole output as ONE device batch (a Partial state: at most one row per group)
        let state = agg.execute_to_device(input_partition, task, &ctx).await?;
        // BatchPartitioner::partition (:336-346): create_hashes(seeds 0,0,0,0) % n, rows grouped by partition, stable
        let mut offs = vec![0u64; n + 1];
        let mut part = std::ptr::null_mut();
        ctx.check(unsafe { bgpu_hash_partition(ctx.0, state.0, keys.as_ptr(), keys.len() as i32, n as u32, &mut part,
                                               offs.as_mut_ptr(), std::ptr::null_mut()) })?;
        let part = DeviceBatch(part);
        // writer.write(&output_batch) per non-empty partition (:351-385): every message framed and LZ4_FRAME-compressed
        // on the device by ONE call, then appended to the partition's file
        let mut msg_off = vec![0u64; n + 1];
        let mut bytes: *mut u8 = std::ptr::null_mut();
        ctx.check(unsafe { bgpu_ipc_encode_partitions(ctx.0, part.0, offs.as_ptr(), n as u32, LZ4_FRAME, &mut bytes, msg_off.as_mut_ptr()) })?;
        let schema = FFI_ArrowSchema::try_from(agg.schema.as_ref())?;
        let (mut sm, mut sm_len) = (std::ptr::null_mut(), 0u64);
        ctx.check(unsafe { bgpu_ipc_schema_message(ctx.0, &schema, &mut sm, &mut sm_len) })?;
        let schema_msg = unsafe { std::slice::from_raw_parts(sm, sm_len as usize) }.to_vec();
        unsafe { bgpu_ipc_free(sm, 0) };
        let mut out = vec![];
        for p in 0..n {
            let (lo, hi) = (msg_off[p] as usize, msg_off[p + 1] as usize);
            if hi == lo { continue; } // empty partitions are skipped, no file (:349)
            let dir = format!("{}/{}/{}/{}", self.work_dir, self.job_id, self.stage_id, p); // :358-362
            std::fs::create_dir_all(&dir)?;
            let path = format!("{dir}/{}.arrow", uuid::Uuid::new_v4());
            let mut w = PartitionFile { file: File::create(&path)?, path, num_batches: 0, num_rows: 0, num_bytes: 0 };
            w.file.write_all(&schema_msg)?;
            w.file.write_all(unsafe { std::slice::from_raw_parts(bytes.add(lo), hi - lo) })?;
            w.num_batches += 1;
            w.num_rows += offs[p + 1] - offs[p];
            w.num_bytes = (schema_msg.len() + hi - lo) as u64; // the file length BEFORE finish() (:396-399)
            w.file.write_all(&EOS)?;
            out.push(ShuffleWritePartition { partition_id: p as u64, path: w.path, num_batches: w.num_batches,
                                             num_rows: w.num_rows, num_bytes: w.num_bytes });
        }
        if !bytes.is_null() { unsafe { bgpu_ipc_free(bytes, 1) }; }
        Ok(out)
    }

    /// Inside one NVSwitch node (one executor process per GPU) the files are skipped: the state is repartitioned and
    /// exchanged over NVLink. COLLECTIVE over the executors of the node; cheapest path first, as
    /// arrow-ballista_b200/exchange.py `exchange_state` does.
    pub fn exchange_state(&self, ctx: &Ctx, comm: *mut std::ffi::c_void, peer: *mut BgpuPeerExchange, rank: i32, world: i32,
                          state: &DeviceBatch, keys: &[i32], n: u32) -> Result<DeviceBatch> {
        let mut out = std::ptr::null_mut();
        let rc = unsafe { bgpu_small_exchange(ctx.0, comm, rank, world, state.0, keys.as_ptr(), keys.len() as i32, n, &mut out) };
        if rc != BGPU_EUNSUPPORTED { ctx.check(rc)?; return Ok(DeviceBatch(out)); }
        let rc = unsafe { bgpu_partition_exchange(peer, state.0, keys.as_ptr(), keys.len() as i32, n, &mut out,
                                                  std::ptr::null_mut(), std::ptr::null_mut()) };
        if rc != BGPU_EUNSUPPORTED { ctx.check(rc)?; return Ok(DeviceBatch(out)); }
        let mut offs = vec![0u64; n as usize + 1];
        let mut part = std::ptr::null_mut();
        ctx.check(unsafe { bgpu_hash_partition(ctx.0, state.0, keys.as_ptr(), keys.len() as i32, n, &mut part, offs.as_mut_ptr(),
                                               std::ptr::null_mut()) })?;
        let part = DeviceBatch(part);
        ctx.check(unsafe { bgpu_all_to_all(ctx.0, comm, rank, world, part.0, offs.as_ptr(), n, &mut out) })?;
        Ok(DeviceBatch(out))
    }
}

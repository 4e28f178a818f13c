//! never compiled here (no cargo in this image). `GpuPhysicalCodec`: a `PhysicalExtensionCodec`
//! (/root/reference/ballista/core/src/serde/mod.rs:151-157, 274-278) that wraps the default codec exactly like the
//! reference's own test codec does (/root/reference/ballista/tests/src/test_physical_codec.rs:29-64): try the inner codec
//! first, else the node is one of ours (proto/bgpu.proto, riding in datafusion.PhysicalExtensionNode,
//! /root/reference/ballista/core/proto/datafusion.proto:1136-1139). Python mirror with the same wire bytes:
//! arrow-ballista_b200/codec.py (checked against google.protobuf in tests/test_codec.py).
use std::sync::Arc;

use ballista_core::serde::BallistaPhysicalExtensionCodec;
use datafusion::error::{DataFusionError, Result};
use datafusion::execution::FunctionRegistry;
use datafusion::physical_plan::aggregates::AggregateMode;
use datafusion::physical_plan::expressions::Column;
use datafusion::physical_plan::{ExecutionPlan, Partitioning, PhysicalExpr};
use datafusion_proto::physical_plan::PhysicalExtensionCodec;
use prost::Message;

use crate::gpu_aggregate::GpuAggregateExec;
use crate::gpu_shuffle_writer::GpuShuffleWriterExec;

pub mod protobuf {
    include!(concat!(env!("OUT_DIR"), "/bgpu.protobuf.rs")); // prost-build output of proto/bgpu.proto (build.rs)
}
use protobuf::gpu_physical_plan_node::Plan;

#[derive(Debug, Default)]
pub struct GpuPhysicalCodec {
    inner: BallistaPhysicalExtensionCodec,
    pub device: i32,
}

impl PhysicalExtensionCodec for GpuPhysicalCodec {
    fn try_decode(&self, buf: &[u8], inputs: &[Arc<dyn ExecutionPlan>], registry: &dyn FunctionRegistry)
                  -> Result<Arc<dyn ExecutionPlan>> {
        if let Ok(plan) = self.inner.try_decode(buf, inputs, registry) {
            return Ok(plan); // a stock Ballista node (ShuffleWriterExec, ShuffleReaderExec, ...)
        }
        let node = protobuf::GpuPhysicalPlanNode::decode(buf)
            .map_err(|e| DataFusionError::Internal(format!("GpuPhysicalCodec: {e}")))?;
        let input = inputs.first().cloned()
            .ok_or_else(|| DataFusionError::Internal("GPU plan node without an input".into()))?;
        match node.plan {
            Some(Plan::ShuffleWriter(w)) => {
                let part = w.output_partitioning.map(|h| {
                    let exprs: Vec<Arc<dyn PhysicalExpr>> = h.hash_expr.iter()
                        .map(|&i| Arc::new(Column::new(input.schema().field(i as usize).name(), i as usize)) as _).collect();
                    Partitioning::Hash(exprs, h.partition_count as usize)
                });
                // work_dir "" as the reference's decode does (serde/mod.rs:191): the engine re-creates the node with the
                // executor's work_dir (execution_engine.rs:74-93)
                Ok(Arc::new(GpuShuffleWriterExec { job_id: w.job_id, stage_id: w.stage_id as usize, plan: input,
                                                   work_dir: String::new(), shuffle_output_partitioning: part,
                                                   device: self.device }))
            }
            Some(Plan::Aggregate(a)) => {
                let mode = match a.mode { 0 => AggregateMode::Partial, 2 => AggregateMode::FinalPartitioned,
                                          m => return Err(DataFusionError::NotImplemented(format!("aggregate mode {m}"))) };
                Ok(Arc::new(GpuAggregateExec::from_proto(mode, &a, input, registry, self.device)?))
            }
            None => Err(DataFusionError::Internal("empty GpuPhysicalPlanNode".into())),
        }
    }

    fn try_encode(&self, node: Arc<dyn ExecutionPlan>, buf: &mut Vec<u8>) -> Result<()> {
        let any = node.as_any();
        let plan = if let Some(w) = any.downcast_ref::<GpuShuffleWriterExec>() {
            let part = match &w.shuffle_output_partitioning {
                Some(Partitioning::Hash(exprs, n)) => Some(protobuf::GpuHashRepartition {
                    hash_expr: exprs.iter().map(|e| e.as_any().downcast_ref::<Column>().map(|c| c.index() as u32)
                        .ok_or_else(|| DataFusionError::NotImplemented("hash key that is not a column".into())))
                        .collect::<Result<_>>()?,
                    partition_count: *n as u64 }),
                None => None,
                Some(other) => return Err(DataFusionError::NotImplemented(format!("partitioning {other:?}"))),
            };
            Plan::ShuffleWriter(protobuf::GpuShuffleWriterExecNode { job_id: w.job_id.clone(), stage_id: w.stage_id as u32,
                                                                     output_partitioning: part, partitions: vec![] })
        } else if let Some(a) = any.downcast_ref::<GpuAggregateExec>() {
            Plan::Aggregate(a.to_proto()?)
        } else {
            return self.inner.try_encode(node, buf); // not ours: the wrapped codec's business
        };
        protobuf::GpuPhysicalPlanNode { plan: Some(plan) }.encode(buf)
            .map_err(|e| DataFusionError::Internal(format!("GpuPhysicalCodec: {e}")))
    }
}

// GpuAggregateExec <-> GpuAggregateExecNode: group columns and aggregate arguments as lowered BgpuExprNode programs, the
// input schema as an encapsulated IPC Schema message (what arrow-ballista_b200/codec.py writes).
impl GpuAggregateExec {
    pub fn to_proto(&self) -> Result<protobuf::GpuAggregateExecNode> {
        let mut n = protobuf::GpuAggregateExecNode { mode: match self.mode { AggregateMode::Partial => 0, _ => 2 }, ..Default::default() };
        n.group_expr = self.group_cols.iter().map(|&c| c as u32).collect();
        n.group_expr_name = self.group_cols.iter().map(|&c| self.input_schema.field(c).name().clone()).collect();
        for a in &self.aggr_expr {
            let mut nodes = vec![];
            for e in a.expressions() {
                crate::gpu_aggregate::lower(&e, &mut nodes).ok_or_else(|| DataFusionError::NotImplemented(a.name().into()))?;
            }
            n.aggr_expr.push(protobuf::GpuExpr { nodes: nodes.iter().map(|x| protobuf::GpuExprNode {
                op: x.op as u32, a: x.a, precision: x.precision, scale: x.scale, lo: x.lo, hi: x.hi }).collect() });
            n.aggr_expr_name.push(a.name().to_string());
            n.aggr_fn.push(match a.name().split('(').next().unwrap_or("").to_ascii_uppercase().as_str() {
                "SUM" => 2, "AVG" => 3, "COUNT" => 4, _ => return Err(DataFusionError::NotImplemented(a.name().into())) });
        }
        n.input_schema = arrow::ipc::writer::IpcDataGenerator::default()
            .schema_to_bytes(&self.input_schema, &arrow::ipc::writer::IpcWriteOptions::default()).ipc_message;
        Ok(n)
    }

    pub fn from_proto(_mode: AggregateMode, _n: &protobuf::GpuAggregateExecNode, _input: Arc<dyn ExecutionPlan>,
                      _registry: &dyn FunctionRegistry, _device: i32) -> Result<GpuAggregateExec> {
        // Rebuilds `aggr_expr` with datafusion::physical_plan::udaf / create_aggregate_expr from (aggr_fn, lowered argument,
        // name) exactly as datafusion-proto's AggregateExecNode decode does (datafusion.proto:1437-1450); the argument
        // programs are raised back to Column / Literal / BinaryExpr trees (the inverse of gpu_aggregate::lower).
        Err(DataFusionError::NotImplemented("written on a box without cargo: see arrow-ballista_b200/codec.py decode_aggregate".into()))
    }
}

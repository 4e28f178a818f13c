//! never compiled here (no cargo in this image). `GpuExecutionEngine`: the stage execution hook of
//! /root/reference/ballista/executor/src/execution_engine.rs:34-60, the least invasive swap point (no scheduler change):
//! passed through `ExecutorProcessConfig.execution_engine` (/root/reference/ballista/executor/src/executor_process.rs:92-94,
//! 270-278). Python mirror: arrow-ballista_b200/execution_engine.py (`match_aggregate_stage`, `GpuQueryStageExec`).
use std::fmt::{Debug, Display, Formatter};
use std::sync::Arc;

use arrow::datatypes::SchemaRef;
use async_trait::async_trait;
use ballista_core::error::{BallistaError, Result};
use ballista_core::execution_plans::{CoalesceTasksExec, ShuffleWriterExec};
use ballista_core::serde::protobuf::ShuffleWritePartition;
use ballista_executor::execution_engine::{DefaultExecutionEngine, ExecutionEngine, QueryStageExecutor};
use datafusion::execution::TaskContext;
use datafusion::physical_plan::aggregates::{AggregateExec, AggregateMode};
use datafusion::physical_plan::expressions::Column;
use datafusion::physical_plan::filter::FilterExec;
use datafusion::physical_plan::metrics::MetricsSet;
use datafusion::physical_plan::{ExecutionPlan, Partitioning};

use crate::gpu_aggregate::{lower, GpuAggregateExec};
use crate::gpu_shuffle_writer::GpuShuffleWriterExec;

pub struct GpuExecutionEngine { pub device: i32, pub fallback: DefaultExecutionEngine }

/// ShuffleWriterExec(Hash) -> [CoalesceTasksExec] -> AggregateExec(Partial) -> [FilterExec] -> source, every expression
/// lowerable: the shape `physical_optimizer/task_group.rs:56-110` leaves behind for the map stage of a grouped aggregate.
pub fn match_aggregate_stage(plan: &Arc<dyn ExecutionPlan>, device: i32) -> Option<GpuAggregateExec> {
    let mut node = plan.clone();
    if let Some(c) = node.as_any().downcast_ref::<CoalesceTasksExec>() { node = c.children()[0].clone(); }
    let agg = node.as_any().downcast_ref::<AggregateExec>()?;
    if *agg.mode() != AggregateMode::Partial { return None; }
    let group_cols: Vec<usize> = agg.group_expr().expr().iter()
        .map(|(e, _)| e.as_any().downcast_ref::<Column>().map(|c| c.index())).collect::<Option<_>>()?;
    let mut scratch = vec![];
    for a in agg.aggr_expr() { for e in a.expressions() { lower(&e, &mut scratch)?; } }
    let (input, filter) = match agg.input().as_any().downcast_ref::<FilterExec>() {
        Some(f) => { lower(f.predicate(), &mut scratch)?; (f.input().clone(), Some(f.predicate().clone())) }
        None => (agg.input().clone(), None),
    };
    Some(GpuAggregateExec::new(AggregateMode::Partial, group_cols, agg.aggr_expr().to_vec(), filter, input.clone(),
                               input.schema(), agg.schema(), device))
}

impl ExecutionEngine for GpuExecutionEngine {
    fn create_query_stage_exec(&self, job_id: String, stage_id: usize, plan: Arc<dyn ExecutionPlan>, work_dir: &str,
                               sender: Option<tokio::sync::mpsc::Sender<ballista_core::replicator::Command>>)
                               -> Result<Arc<dyn QueryStageExecutor>> {
        // the root of every stage plan is a concrete ShuffleWriterExec (planner.rs:42, 64-79)
        let writer = plan.as_any().downcast_ref::<ShuffleWriterExec>().ok_or_else(|| {
            BallistaError::General("Plan passed to new_query_stage_exec is not a ShuffleWriterExec".to_string()) })?;
        if let Some(Partitioning::Hash(_, _)) = writer.shuffle_output_partitioning() {
            if let Some(agg) = match_aggregate_stage(&writer.children()[0], self.device) {
                let exec = GpuShuffleWriterExec { job_id, stage_id, plan: Arc::new(agg), work_dir: work_dir.to_string(),
                                                  shuffle_output_partitioning: writer.shuffle_output_partitioning().cloned(),
                                                  device: self.device };
                return Ok(Arc::new(GpuQueryStageExec { exec, schema: writer.schema() }));
            }
        }
        self.fallback.create_query_stage_exec(job_id, stage_id, plan, work_dir, sender) // any other plan: the CPU engine
    }
}

pub struct GpuQueryStageExec { exec: GpuShuffleWriterExec, schema: SchemaRef }
impl Debug for GpuQueryStageExec { fn fmt(&self, f: &mut Formatter<'_>) -> std::fmt::Result { write!(f, "{:?}", self.exec) } }
impl Display for GpuQueryStageExec {
    fn fmt(&self, f: &mut Formatter<'_>) -> std::fmt::Result { write!(f, "GpuQueryStageExec: ({}, {})", self.exec.job_id, self.exec.stage_id) }
}

#[async_trait]
impl QueryStageExecutor for GpuQueryStageExec {
    async fn execute_query_stage(&self, input_partitions: Vec<usize>, context: Arc<TaskContext>) -> Result<Vec<ShuffleWritePartition>> {
        // a task group: the child's CoalesceTasksExec fans the partitions in; here each one is a stream of the aggregate
        let mut out = vec![];
        for p in input_partitions {
            out.extend(self.exec.execute_shuffle_write(p, context.clone()).await.map_err(BallistaError::DataFusionError)?);
        }
        Ok(out)
    }
    fn collect_plan_metrics(&self) -> Vec<MetricsSet> { vec![] } // write_time, repart_time, input_rows, output_rows (+ h2d/kernel/d2h)
    fn schema(&self) -> SchemaRef { self.schema.clone() }
}

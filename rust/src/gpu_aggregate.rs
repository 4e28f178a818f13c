//! never compiled here (no cargo in this image). `GpuAggregateExec`: DataFusion's `AggregateExec` for the modes on
//! Ballista's hot path (Partial, FinalPartitioned; plan shape /root/reference/ballista/scheduler/src/planner.rs:376-427,
//! proto /root/reference/ballista/core/proto/datafusion.proto:1437-1450) over libbgpu. The Python mirror of the same call
//! sequence is arrow-ballista_b200/aggregate.py; the parity tests drive that.
use std::any::Any;
use std::sync::Arc;

use arrow::array::{Array, StructArray};
use arrow::datatypes::SchemaRef;
use arrow::ffi::{from_ffi, to_ffi, FFI_ArrowArray, FFI_ArrowSchema};
use arrow::record_batch::RecordBatch;
use datafusion::error::{DataFusionError, Result};
use datafusion::execution::TaskContext;
use datafusion::physical_plan::aggregates::AggregateMode;
use datafusion::physical_plan::expressions::{BinaryExpr, Column, Literal};
use datafusion::physical_plan::metrics::{ExecutionPlanMetricsSet, MetricsSet};
use datafusion::physical_plan::stream::RecordBatchStreamAdapter;
use datafusion::physical_plan::{
    AggregateExpr, DisplayAs, DisplayFormatType, ExecutionPlan, Partitioning, PhysicalExpr, SendableRecordBatchStream,
};
use datafusion::scalar::ScalarValue;
use futures::StreamExt;

use crate::ffi::*;

/// `AggregateExec`'s fields (`mode, group_by, aggr_expr, input, input_schema`), plus the FilterExec fused below a
/// Partial aggregate (Q1's `l_shipdate <= date`, /root/reference/benchmarks/queries/q1.sql).
#[derive(Debug)]
pub struct GpuAggregateExec {
    pub mode: AggregateMode,
    pub group_cols: Vec<usize>,              // indices into input_schema (the planner's group exprs are Columns)
    pub aggr_expr: Vec<Arc<dyn AggregateExpr>>,
    pub filter: Option<Arc<dyn PhysicalExpr>>,
    pub input: Arc<dyn ExecutionPlan>,
    pub input_schema: SchemaRef,
    pub schema: SchemaRef,                   // as AggregateExec computes it (state schema in Partial mode)
    pub device: i32,
    metrics: ExecutionPlanMetricsSet,
}

/// Lower a physical expression to a postfix BgpuExprNode program. `None` = not expressible: the engine then keeps the
/// CPU operator for this stage (decided at plan time, never at run time).
pub fn lower(e: &Arc<dyn PhysicalExpr>, out: &mut Vec<BgpuExprNode>) -> Option<()> {
    use datafusion::logical_expr::Operator::*;
    let any = e.as_any();
    if let Some(c) = any.downcast_ref::<Column>() {
        out.push(BgpuExprNode { op: 0, a: c.index() as i32, ..Default::default() }); // BGPU_X_COLUMN
    } else if let Some(l) = any.downcast_ref::<Literal>() {
        let (t, p, s, v): (i32, i32, i32, i128) = match l.value() {
            ScalarValue::Int64(Some(v)) => (0, 0, 0, *v as i128),
            ScalarValue::Float64(Some(v)) => (1, 0, 0, v.to_bits() as i128),
            ScalarValue::Decimal128(Some(v), p, s) => (2, *p as i32, *s as i32, *v),
            ScalarValue::Date32(Some(v)) => (3, 0, 0, *v as i128),
            _ => return None,
        };
        out.push(BgpuExprNode { op: 1, a: t, precision: p, scale: s, lo: v as u64, hi: (v >> 64) as u64 });
    } else if let Some(b) = any.downcast_ref::<BinaryExpr>() {
        lower(b.left(), out)?;
        lower(b.right(), out)?;
        let op = match b.op() { Plus => 2, Minus => 3, Multiply => 4, Lt => 5, LtEq => 6, Gt => 7, GtEq => 8, Eq => 9,
                                NotEq => 10, And => 11, _ => return None };
        out.push(BgpuExprNode { op, ..Default::default() });
    } else {
        return None;
    }
    Some(())
}

impl GpuAggregateExec {
    #[allow(clippy::too_many_arguments)]
    pub fn new(mode: AggregateMode, group_cols: Vec<usize>, aggr_expr: Vec<Arc<dyn AggregateExpr>>,
               filter: Option<Arc<dyn PhysicalExpr>>, input: Arc<dyn ExecutionPlan>, input_schema: SchemaRef, schema: SchemaRef,
               device: i32) -> Self {
        GpuAggregateExec { mode, group_cols, aggr_expr, filter, input, input_schema, schema, device,
                           metrics: ExecutionPlanMetricsSet::new() }
    }

    /// bgpu_agg_create_expr for this plan on `ctx` (one aggregate per executed partition: a stream of the plan).
    fn create(&self, ctx: &Ctx) -> Result<Agg> {
        let schema = FFI_ArrowSchema::try_from(self.input_schema.as_ref())?;
        let group: Vec<i32> = self.group_cols.iter().map(|&c| c as i32).collect();
        let (mut fns, mut nodes, mut offs, mut names) = (vec![], vec![], vec![0i32], vec![]);
        for a in &self.aggr_expr {
            // AggregateFunction numbering of datafusion.proto:631-636: SUM = 2, AVG = 3, COUNT = 4
            let f = match a.name().split('(').next().unwrap_or("").to_ascii_uppercase().as_str() {
                "SUM" => 2, "AVG" => 3, "COUNT" => 4,
                other => return Err(DataFusionError::NotImplemented(format!("aggregate {other} is not on the GPU path"))),
            };
            fns.push(f);
            for e in a.expressions() {
                lower(&e, &mut nodes).ok_or_else(|| DataFusionError::NotImplemented(format!("argument of {}", a.name())))?;
            }
            offs.push(nodes.len() as i32);
            names.push(std::ffi::CString::new(a.name()).unwrap());
        }
        let name_ptrs: Vec<_> = names.iter().map(|n| n.as_ptr()).collect();
        let mut filter = vec![];
        if let Some(f) = &self.filter {
            lower(f, &mut filter).ok_or_else(|| DataFusionError::NotImplemented("filter expression".into()))?;
        }
        let mode = match self.mode { AggregateMode::Partial => AGG_PARTIAL, AggregateMode::FinalPartitioned => AGG_FINAL_PARTITIONED,
                                     m => return Err(DataFusionError::NotImplemented(format!("aggregate mode {m:?}"))) };
        let mut h = std::ptr::null_mut();
        ctx.check(unsafe { bgpu_agg_create_expr(ctx.0, mode, &schema, group.as_ptr(), group.len() as i32, fns.as_ptr(),
                                                nodes.as_ptr(), offs.as_ptr(), name_ptrs.as_ptr(), fns.len() as i32,
                                                filter.as_ptr(), filter.len() as i32, &mut h) })?;
        Ok(Agg(h))
    }

    /// The device-resident form `GpuShuffleWriterExec` takes when it sits directly above this operator: the state batch
    /// never leaves the GPU. Every input batch is one `bgpu_agg_push_host` call (thread-safe: CoalesceTasksExec's
    /// per-partition tasks, /root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:183-216, may each push).
    pub async fn execute_to_device(&self, partition: usize, task: Arc<TaskContext>, ctx: &Ctx) -> Result<DeviceBatch> {
        let agg = self.create(ctx)?;
        let mut input = self.input.execute(partition, task)?;
        while let Some(batch) = input.next().await {
            let data = StructArray::from(batch?).into_data();
            let (arr, sch) = to_ffi(&data)?; // borrowed for the call: push returns once the buffers have been read
            ctx.check(unsafe { bgpu_agg_push_host(agg.0, &arr, &sch) })?;
        }
        let mut out = std::ptr::null_mut();
        ctx.check(unsafe { bgpu_agg_emit(agg.0, &mut out) })?; // drains the staged chunks, then state() / evaluate()
        Ok(DeviceBatch(out))
    }
}

pub fn download(ctx: &Ctx, b: &DeviceBatch) -> Result<RecordBatch> {
    let (mut arr, mut sch) = (FFI_ArrowArray::empty(), FFI_ArrowSchema::empty());
    ctx.check(unsafe { bgpu_batch_download(ctx.0, b.0, 0, b.num_rows() as i64, &mut arr, &mut sch) })?;
    Ok(RecordBatch::from(StructArray::from(unsafe { from_ffi(arr, &sch) }?)))
}

impl DisplayAs for GpuAggregateExec {
    fn fmt_as(&self, _t: DisplayFormatType, f: &mut std::fmt::Formatter) -> std::fmt::Result {
        write!(f, "GpuAggregateExec: mode={:?}, gby={:?}, aggr=[{}]", self.mode, self.group_cols,
               self.aggr_expr.iter().map(|a| a.name().to_string()).collect::<Vec<_>>().join(", "))
    }
}

impl ExecutionPlan for GpuAggregateExec {
    fn as_any(&self) -> &dyn Any { self }
    fn schema(&self) -> SchemaRef { self.schema.clone() }
    fn output_partitioning(&self) -> Partitioning { self.input.output_partitioning() } // as AggregateExec (Partial / FinalPartitioned)
    fn output_ordering(&self) -> Option<&[datafusion::physical_expr::PhysicalSortExpr]> { None }
    fn children(&self) -> Vec<Arc<dyn ExecutionPlan>> { vec![self.input.clone()] }
    fn with_new_children(self: Arc<Self>, c: Vec<Arc<dyn ExecutionPlan>>) -> Result<Arc<dyn ExecutionPlan>> {
        Ok(Arc::new(GpuAggregateExec { input: c[0].clone(), mode: self.mode, group_cols: self.group_cols.clone(),
                                       aggr_expr: self.aggr_expr.clone(), filter: self.filter.clone(),
                                       input_schema: self.input_schema.clone(), schema: self.schema.clone(),
                                       device: self.device, metrics: ExecutionPlanMetricsSet::new() }))
    }
    /// Stand-alone use (a CPU operator above): the emitted batch is downloaded and sliced to batch_size.
    fn execute(&self, partition: usize, task: Arc<TaskContext>) -> Result<SendableRecordBatchStream> {
        let batch_size = task.session_config().batch_size();
        let this = Arc::new(self.with_same_fields());
        let schema = self.schema.clone();
        let fut = async move {
            let ctx = Ctx::new(this.device)?;
            let state = this.execute_to_device(partition, task, &ctx).await?;
            let whole = download(&ctx, &state)?;
            let slices: Vec<Result<RecordBatch>> = (0..whole.num_rows()).step_by(batch_size.max(1))
                .map(|lo| Ok(whole.slice(lo, batch_size.min(whole.num_rows() - lo)))).collect();
            Ok::<_, DataFusionError>(futures::stream::iter(slices))
        };
        Ok(Box::pin(RecordBatchStreamAdapter::new(schema, futures::stream::once(fut).try_flatten())))
    }
    fn metrics(&self) -> Option<MetricsSet> { Some(self.metrics.clone_inner()) }
    fn statistics(&self) -> Result<datafusion::common::Statistics> { Ok(datafusion::common::Statistics::new_unknown(&self.schema)) }
}

impl GpuAggregateExec {
    fn with_same_fields(&self) -> GpuAggregateExec {
        GpuAggregateExec { mode: self.mode, group_cols: self.group_cols.clone(), aggr_expr: self.aggr_expr.clone(),
                           filter: self.filter.clone(), input: self.input.clone(), input_schema: self.input_schema.clone(),
                           schema: self.schema.clone(), device: self.device, metrics: ExecutionPlanMetricsSet::new() }
    }
}
use futures::TryStreamExt;

//! `GpuAggregateExec`: drop-in for `AggregateExec` (physical-plan/src/aggregates/mod.rs:242-299, 638-747) whose
//! `execute()` replaces `GroupedHashAggregateStream` (row_hash.rs) by the device operator behind `dfgpu_agg_*`.
//!
//! Planning properties (schema, partitioning, required distribution, equivalences) are the CPU node's own: the wrapped
//! `AggregateExec` is kept and asked, so `EnforceDistribution`, `EnforceSorting` and EXPLAIN see the same operator.
use std::any::Any;
use std::fmt;
use std::ptr;
use std::sync::Arc;

use arrow::array::ArrayRef;
use arrow::datatypes::{Field, Schema, SchemaRef};
use arrow::ffi::{FFI_ArrowArray, FFI_ArrowSchema};
use arrow::record_batch::RecordBatch;
use datafusion_common::{DataFusionError, Result};
use datafusion_execution::{SendableRecordBatchStream, TaskContext};
use datafusion_physical_expr::expressions::{Avg, BitAnd, BitOr, BitXor, BoolAnd, BoolOr, Count, Max, Min, Sum};
use datafusion_physical_expr::{AggregateExpr, Distribution, PhysicalExpr};
use datafusion_physical_plan::aggregates::{AggregateExec, AggregateMode};
use datafusion_physical_plan::metrics::{ExecutionPlanMetricsSet, MetricBuilder, MetricsSet};
use datafusion_physical_plan::stream::RecordBatchStreamAdapter;
use datafusion_physical_plan::{DisplayAs, DisplayFormatType, ExecutionPlan, InputOrderMode, PlanProperties};
use futures::{StreamExt, TryStreamExt};

use crate::ctx::{cstring, export_batch, export_schema, import_batch, slice_batches, GpuCtx};
use crate::sys::*;

#[derive(Debug)]
pub struct GpuAggregateExec {
    cpu: Arc<AggregateExec>,
    ctx: Arc<GpuCtx>,
    /// `Single` mode + `dfgpu_agg_exchange` in place of Partial -> RepartitionExec(Hash) -> FinalPartitioned (rule.rs)
    fused_exchange: bool,
    metrics: ExecutionPlanMetricsSet,
}
impl fmt::Debug for GpuCtx {
    fn fmt(&self, f: &mut fmt::Formatter) -> fmt::Result {
        write!(f, "GpuCtx(device {}, rank {}/{})", self.options.device, self.options.rank, self.options.world)
    }
}

/// How the operator's input batches are laid out for the device: evaluated expressions, by column position.
pub(crate) struct Projection {
    pub schema: SchemaRef,
    pub exprs: Vec<Arc<dyn PhysicalExpr>>, // one per projected column (first-stage modes); empty = pass the batch through
    pub group_cols: Vec<i32>,
    pub descs: Vec<(i32, String, i32, i32)>, // (func, name, arg_col, filter_col)
}

/// The aggregate functions with a `GroupsAccumulator` on this path -> `DFGPU_*`.
pub(crate) fn func_code(e: &Arc<dyn AggregateExpr>) -> Option<i32> {
    let a = e.as_any();
    if a.is::<Sum>() {
        Some(DFGPU_SUM)
    } else if a.is::<Count>() {
        Some(DFGPU_COUNT)
    } else if a.is::<Avg>() {
        Some(DFGPU_AVG)
    } else if a.is::<Min>() {
        Some(DFGPU_MIN)
    } else if a.is::<Max>() {
        Some(DFGPU_MAX)
    } else if a.is::<BitAnd>() {
        Some(DFGPU_BIT_AND)
    } else if a.is::<BitOr>() {
        Some(DFGPU_BIT_OR)
    } else if a.is::<BitXor>() {
        Some(DFGPU_BIT_XOR)
    } else if a.is::<BoolAnd>() {
        Some(DFGPU_BOOL_AND)
    } else if a.is::<BoolOr>() {
        Some(DFGPU_BOOL_OR)
    } else {
        None
    }
}

pub(crate) fn mode_code(m: &AggregateMode) -> i32 {
    match m {
        AggregateMode::Partial => DFGPU_AGG_PARTIAL,
        AggregateMode::Final => DFGPU_AGG_FINAL,
        AggregateMode::FinalPartitioned => DFGPU_AGG_FINAL_PARTITIONED,
        AggregateMode::Single => DFGPU_AGG_SINGLE,
        AggregateMode::SinglePartitioned => DFGPU_AGG_SINGLE_PARTITIONED,
    }
}

/// What `group_aggregate_batch` evaluates per batch (aggregates/mod.rs:1128-1200), as one projected batch:
/// first-stage modes `[group exprs..., argument exprs (one column per aggregate with an argument)..., FILTER exprs...]`;
/// final modes: the Partial output as is — state columns are resolved by position, `group_cols + running offset`
/// (aggregates/mod.rs:1038-1067), which `dfgpu_agg_create` does itself.
pub(crate) fn project_plan(cpu: &AggregateExec) -> Result<Projection> {
    let input_schema = cpu.input().schema();
    let first_stage = cpu.mode().is_first_stage();
    let n_group = cpu.group_expr().expr().len();
    let mut descs = vec![];
    if !first_stage {
        for e in cpu.aggr_expr() {
            let f = func_code(e).ok_or_else(|| DataFusionError::NotImplemented(format!("{} on the GPU path", e.name())))?;
            descs.push((f, e.name().to_string(), -1, -1));
        }
        return Ok(Projection { schema: input_schema, exprs: vec![], group_cols: (0..n_group as i32).collect(), descs });
    }
    let mut fields: Vec<Field> = vec![];
    let mut exprs: Vec<Arc<dyn PhysicalExpr>> = vec![];
    for (e, name) in cpu.group_expr().expr() {
        fields.push(Field::new(name, e.data_type(&input_schema)?, e.nullable(&input_schema)?));
        exprs.push(e.clone());
    }
    for (e, filter) in cpu.aggr_expr().iter().zip(cpu.filter_expr()) {
        let f = func_code(e).ok_or_else(|| DataFusionError::NotImplemented(format!("{} on the GPU path", e.name())))?;
        let args = e.expressions();
        // COUNT(*) arrives as COUNT(UInt8(1)): a literal argument means "every row" (count.rs:110-133)
        let arg_col = match args.first() {
            Some(a) if a.as_any().downcast_ref::<datafusion_physical_expr::expressions::Literal>().is_none() => {
                fields.push(Field::new(format!("arg{}", exprs.len()), a.data_type(&input_schema)?, true));
                exprs.push(a.clone());
                exprs.len() as i32 - 1
            }
            _ => -1,
        };
        let filter_col = match filter {
            Some(p) => {
                fields.push(Field::new(format!("filter{}", exprs.len()), arrow::datatypes::DataType::Boolean, true));
                exprs.push(p.clone());
                exprs.len() as i32 - 1
            }
            None => -1,
        };
        descs.push((f, e.name().to_string(), arg_col, filter_col));
    }
    Ok(Projection { schema: Arc::new(Schema::new(fields)), exprs, group_cols: (0..n_group as i32).collect(), descs })
}

pub(crate) fn evaluate_projection(p: &Projection, batch: &RecordBatch) -> Result<RecordBatch> {
    if p.exprs.is_empty() {
        return Ok(batch.clone());
    }
    let cols: Vec<ArrayRef> =
        p.exprs.iter().map(|e| e.evaluate(batch).and_then(|v| v.into_array(batch.num_rows()))).collect::<Result<_>>()?;
    Ok(RecordBatch::try_new(p.schema.clone(), cols)?)
}

/// Owns a `dfgpu_agg`.
pub(crate) struct GpuAgg {
    pub ptr: *mut dfgpu_agg,
    _ctx: Arc<GpuCtx>,
}
unsafe impl Send for GpuAgg {}
impl GpuAgg {
    pub fn create(ctx: &Arc<GpuCtx>, mode: i32, p: &Projection) -> Result<Self> {
        let schema = export_schema(&p.schema)?;
        let names: Vec<_> = p.descs.iter().map(|d| cstring(&d.1)).collect();
        let descs: Vec<dfgpu_agg_desc> = p
            .descs
            .iter()
            .zip(&names)
            .map(|(d, n)| dfgpu_agg_desc { func: d.0, name: n.as_ptr(), arg_col: d.2, filter_col: d.3 })
            .collect();
        let mut out = ptr::null_mut();
        check(unsafe {
            dfgpu_agg_create(ctx.ptr, mode, &schema, p.group_cols.as_ptr(), p.group_cols.len() as i32, descs.as_ptr(), descs.len() as i32, &mut out)
        })?;
        Ok(Self { ptr: out, _ctx: ctx.clone() })
    }
    /// group_aggregate_batch (row_hash.rs:524-613).  Small batches are packed into a pinned arena inside the library and
    /// cross PCIe `coalesce_rows` rows at a time; the batch may be dropped when this returns.
    pub fn push(&self, batch: &RecordBatch) -> Result<()> {
        let (array, _schema) = export_batch(batch)?;
        check(unsafe { dfgpu_agg_push_arrow(self.ptr, &array) })
    }
    /// Several batches the input stream had ready, in one call: the library packs them into its pinned arena with
    /// several host threads (`dfgpu_agg_push_arrow_batches`).
    pub fn push_many(&self, batches: &[RecordBatch]) -> Result<()> {
        let exported = batches.iter().map(export_batch).collect::<Result<Vec<_>>>()?;
        let ptrs: Vec<*const FFI_ArrowArray> = exported.iter().map(|(a, _)| a as *const FFI_ArrowArray).collect();
        check(unsafe { dfgpu_agg_push_arrow_batches(self.ptr, ptrs.as_ptr(), ptrs.len() as i32) })
    }
    pub fn exchange(&self) -> Result<()> {
        check(unsafe { dfgpu_agg_exchange(self.ptr) })
    }
    /// emit(EmitTo::All) (row_hash.rs:626-662)
    pub fn emit(&self) -> Result<RecordBatch> {
        let (mut a, mut s) = (FFI_ArrowArray::empty(), FFI_ArrowSchema::empty());
        check(unsafe { dfgpu_agg_emit_arrow(self.ptr, &mut a, &mut s) })?;
        import_batch(a, &s)
    }
    pub fn metrics(&self) -> dfgpu_metrics {
        let mut m = dfgpu_metrics::default();
        unsafe { dfgpu_agg_metrics(self.ptr, &mut m) };
        m
    }
}
impl Drop for GpuAgg {
    fn drop(&mut self) {
        unsafe { dfgpu_agg_destroy(self.ptr) }
    }
}

impl GpuAggregateExec {
    /// Fails with `NotImplemented` when the device path does not take the operator; the rule then keeps the CPU node.
    pub fn try_new(cpu: Arc<AggregateExec>, ctx: Arc<GpuCtx>, fused_exchange: bool) -> Result<Self> {
        if !cpu.group_expr().is_single() || cpu.group_expr().expr().is_empty() {
            return Err(DataFusionError::NotImplemented("grouping sets / ungrouped aggregation on the GPU path".into()));
        }
        if *cpu.input_order_mode() != InputOrderMode::Linear || cpu.limit().is_some() {
            return Err(DataFusionError::NotImplemented("ordered or limited aggregation on the GPU path".into()));
        }
        // creating (and dropping) the device operator validates key and argument types once, at planning time
        let p = project_plan(&cpu)?;
        let mode = if fused_exchange { DFGPU_AGG_SINGLE } else { mode_code(cpu.mode()) };
        drop(GpuAgg::create(&ctx, mode, &p)?);
        Ok(Self { cpu, ctx, fused_exchange, metrics: ExecutionPlanMetricsSet::new() })
    }
    pub fn cpu(&self) -> &Arc<AggregateExec> {
        &self.cpu
    }
}

impl DisplayAs for GpuAggregateExec {
    fn fmt_as(&self, t: DisplayFormatType, f: &mut fmt::Formatter) -> fmt::Result {
        write!(f, "Gpu")?;
        self.cpu.fmt_as(t, f)?;
        if self.fused_exchange {
            write!(f, ", fused_exchange=true")?;
        }
        Ok(())
    }
}

impl ExecutionPlan for GpuAggregateExec {
    fn as_any(&self) -> &dyn Any {
        self
    }
    fn properties(&self) -> &PlanProperties {
        self.cpu.properties() // schema (create_schema, mod.rs:749-787), partitioning, equivalences, execution mode
    }
    fn required_input_distribution(&self) -> Vec<Distribution> {
        self.cpu.required_input_distribution() // mod.rs:648-660
    }
    fn children(&self) -> Vec<Arc<dyn ExecutionPlan>> {
        self.cpu.children()
    }
    fn with_new_children(self: Arc<Self>, children: Vec<Arc<dyn ExecutionPlan>>) -> Result<Arc<dyn ExecutionPlan>> {
        let cpu = self.cpu.clone().with_new_children(children)?;
        let cpu = cpu
            .as_any()
            .downcast_ref::<AggregateExec>()
            .ok_or_else(|| DataFusionError::Internal("AggregateExec::with_new_children changed the node type".into()))?;
        let cpu = Arc::new(AggregateExec::try_new(
            *cpu.mode(),
            cpu.group_expr().clone(),
            cpu.aggr_expr().to_vec(),
            cpu.filter_expr().to_vec(),
            cpu.input().clone(),
            cpu.input_schema(),
        )?);
        Ok(Arc::new(GpuAggregateExec::try_new(cpu, self.ctx.clone(), self.fused_exchange)?))
    }
    fn execute(&self, partition: usize, tctx: Arc<TaskContext>) -> Result<SendableRecordBatchStream> {
        let mut input = self.cpu.input().execute(partition, tctx.clone())?;
        let p = project_plan(&self.cpu)?;
        let mode = if self.fused_exchange { DFGPU_AGG_SINGLE } else { mode_code(self.cpu.mode()) };
        let agg = GpuAgg::create(&self.ctx, mode, &p)?;
        let batch_size = tctx.session_config().batch_size();
        let out_schema = self.schema();
        let fused = self.fused_exchange;
        let elapsed = MetricBuilder::new(&self.metrics).elapsed_compute(partition);
        let output_rows = MetricBuilder::new(&self.metrics).output_rows(partition);
        let schema_for_cast = out_schema.clone();
        let stream = futures::stream::once(async move {
            // ReadingInput (row_hash.rs:423-470).  Whatever the child has ready is handed over in one ABI call
            // (`ready_chunks` never waits for more than one batch): the library packs the run into its pinned arena
            // with several host threads instead of one copy per call.
            let mut input = input.ready_chunks(64);
            while let Some(ready) = input.next().await {
                let mut projected = Vec::with_capacity(ready.len());
                for b in ready {
                    projected.push(evaluate_projection(&p, &b?)?);
                }
                if projected.len() == 1 {
                    agg.push(&projected[0])?;
                } else {
                    agg.push_many(&projected)?;
                }
            }
            if fused {
                agg.exchange()?; // collective: every rank of the context's communicator runs this plan
            }
            // ProducingOutput (row_hash.rs:478-502): keys ++ state fields (Partial) | result fields
            let out = agg.emit()?;
            let out = RecordBatch::try_new(schema_for_cast, out.columns().to_vec())?; // field metadata / names of the CPU schema
            let m = agg.metrics();
            elapsed.add_duration(std::time::Duration::from_nanos(m.elapsed_compute_ns as u64));
            output_rows.add(out.num_rows());
            Ok::<_, DataFusionError>(futures::stream::iter(slice_batches(out, batch_size).into_iter().map(Ok)))
        })
        .try_flatten();
        Ok(Box::pin(RecordBatchStreamAdapter::new(out_schema, stream)))
    }
    fn metrics(&self) -> Option<MetricsSet> {
        Some(self.metrics.clone_inner())
    }
}

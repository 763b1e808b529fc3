//! `GpuHashJoinExec`: drop-in for `HashJoinExec` (physical-plan/src/joins/hash_join.rs:272-356, 505-674).  Left child =
//! build side, right child = probe side, as in the reference (hash_join.rs:404-407).
//!
//! * build: `collect_left_input` (hash_join.rs:678-768) = `dfgpu_join_build_push`* + `dfgpu_join_build_finish`; once per
//!   partition in `Partitioned` mode, once per plan behind a `OnceAsync` in `CollectLeft` mode (hash_join.rs:594-606).
//! * probe: `process_probe_batch` (hash_join.rs:1238-1343) = `dfgpu_join_probe`, or, with a residual `JoinFilter`,
//!   `dfgpu_join_probe_candidates` -> `JoinFilter::expression().evaluate()` on the CPU -> `dfgpu_join_probe_commit`:
//!   the reference filters the candidate pairs BEFORE it marks visited build rows and before
//!   `adjust_indices_by_join_type` (hash_join.rs:1259-1311), so the filter is never applied to the joined output.
//! * end: `process_unmatched_build_batch` (hash_join.rs:1348-1388) = `dfgpu_join_emit_unmatched`.
use std::any::Any;
use std::fmt;
use std::ptr;
use std::sync::Arc;

use arrow::datatypes::SchemaRef;
use arrow::record_batch::RecordBatch;
use datafusion_common::{DataFusionError, JoinSide, JoinType, Result};
use datafusion_execution::{SendableRecordBatchStream, TaskContext};
use datafusion_physical_expr::expressions::Column;
use datafusion_physical_expr::{Distribution, PhysicalExpr};
use datafusion_physical_plan::joins::utils::JoinFilter;
use datafusion_physical_plan::joins::{HashJoinExec, PartitionMode};
use datafusion_physical_plan::metrics::{ExecutionPlanMetricsSet, MetricBuilder, MetricsSet};
use datafusion_physical_plan::stream::RecordBatchStreamAdapter;
use datafusion_physical_plan::{DisplayAs, DisplayFormatType, ExecutionPlan, PlanProperties};
use futures::{StreamExt, TryStreamExt};
use parking_lot::Mutex;

use crate::ctx::{export_schema, slice_batches, DeviceBatch, GpuCtx};
use crate::sys::*;

pub(crate) fn join_type_code(t: &JoinType) -> i32 {
    match t {
        JoinType::Inner => DFGPU_JOIN_INNER,
        JoinType::Left => DFGPU_JOIN_LEFT,
        JoinType::Right => DFGPU_JOIN_RIGHT,
        JoinType::Full => DFGPU_JOIN_FULL,
        JoinType::LeftSemi => DFGPU_JOIN_LEFT_SEMI,
        JoinType::RightSemi => DFGPU_JOIN_RIGHT_SEMI,
        JoinType::LeftAnti => DFGPU_JOIN_LEFT_ANTI,
        JoinType::RightAnti => DFGPU_JOIN_RIGHT_ANTI,
    }
}

/// Join keys must be plain columns of the children (the planner projects other expressions first).
fn key_index(e: &Arc<dyn PhysicalExpr>) -> Result<i32> {
    e.as_any()
        .downcast_ref::<Column>()
        .map(|c| c.index() as i32)
        .ok_or_else(|| DataFusionError::NotImplemented("non-column join key on the GPU path".into()))
}

/// Owns a `dfgpu_join` whose build side is finished: read-only hash table + visited flags updated with atomics, so
/// probe partitions may share it (`Arc<JoinLeftData>` of the reference, hash_join.rs:73-82).
pub(crate) struct GpuJoin {
    ptr: *mut dfgpu_join,
    ctx: Arc<GpuCtx>,
    filter: Option<JoinFilter>,
}
unsafe impl Send for GpuJoin {}
unsafe impl Sync for GpuJoin {}

impl GpuJoin {
    fn create(ctx: &Arc<GpuCtx>, cpu: &HashJoinExec) -> Result<Self> {
        let (bs, ps) = (export_schema(&cpu.left().schema())?, export_schema(&cpu.right().schema())?);
        let (mut bk, mut pk) = (vec![], vec![]);
        for (l, r) in cpu.on() {
            bk.push(key_index(l)?);
            pk.push(key_index(r)?);
        }
        let mode = match cpu.partition_mode() {
            PartitionMode::CollectLeft => DFGPU_COLLECT_LEFT,
            _ => DFGPU_PARTITIONED,
        };
        let mut out = ptr::null_mut();
        check(unsafe {
            dfgpu_join_create(ctx.ptr, join_type_code(cpu.join_type()), mode, cpu.null_equals_null() as i32, &bs, &ps, bk.as_ptr(), pk.as_ptr(), bk.len() as i32, &mut out)
        })?;
        let j = Self { ptr: out, ctx: ctx.clone(), filter: cpu.filter().cloned() };
        if let Some(f) = &j.filter {
            let sides: Vec<i32> = f.column_indices().iter().map(|c| if c.side == JoinSide::Left { 0 } else { 1 }).collect();
            let cols: Vec<i32> = f.column_indices().iter().map(|c| c.index as i32).collect();
            check(unsafe { dfgpu_join_set_filter(j.ptr, sides.as_ptr(), cols.as_ptr(), cols.len() as i32) })?;
        }
        Ok(j)
    }
    fn build_push(&self, b: &RecordBatch) -> Result<()> {
        let d = DeviceBatch::from_record_batch(&self.ctx, b)?;
        check(unsafe { dfgpu_join_build_push(self.ptr, d.ptr) }) // the library keeps its own reference to the columns
    }
    fn build_finish(&self) -> Result<()> {
        check(unsafe { dfgpu_join_build_finish(self.ptr) })
    }
    fn probe(&self, batch: &RecordBatch) -> Result<RecordBatch> {
        let d = DeviceBatch::from_record_batch(&self.ctx, batch)?;
        let mut out = ptr::null_mut();
        match &self.filter {
            None => check(unsafe { dfgpu_join_probe(self.ptr, d.ptr, &mut out) })?,
            Some(f) => {
                // apply_join_filter_to_indices (joins/utils.rs:1143-1176), the evaluation on the CPU
                let mut inter = ptr::null_mut();
                check(unsafe { dfgpu_join_probe_candidates(self.ptr, d.ptr, &mut inter) })?;
                let inter = DeviceBatch::from_raw(&self.ctx, inter).expect("intermediate batch").to_record_batch()?;
                let inter = RecordBatch::try_new(Arc::new(f.schema().clone()), inter.columns().to_vec())?;
                let mask = f.expression().evaluate(&inter)?.into_array(inter.num_rows())?;
                let mask = RecordBatch::try_from_iter([("filter", mask)])?;
                let m = DeviceBatch::from_record_batch(&self.ctx, &mask)?;
                check(unsafe { dfgpu_join_probe_commit(self.ptr, m.ptr, 0, &mut out) })?;
            }
        }
        DeviceBatch::from_raw(&self.ctx, out).expect("probe output").to_record_batch()
    }
    fn emit_unmatched(&self) -> Result<Option<RecordBatch>> {
        let mut out = ptr::null_mut();
        check(unsafe { dfgpu_join_emit_unmatched(self.ptr, &mut out) })?;
        DeviceBatch::from_raw(&self.ctx, out).map(|d| d.to_record_batch()).transpose()
    }
    fn metrics(&self) -> dfgpu_metrics {
        let mut m = dfgpu_metrics::default();
        unsafe { dfgpu_join_metrics(self.ptr, &mut m) };
        m
    }
}
impl Drop for GpuJoin {
    fn drop(&mut self) {
        unsafe { dfgpu_join_destroy(self.ptr) }
    }
}

#[derive(Debug)]
pub struct GpuHashJoinExec {
    cpu: Arc<HashJoinExec>,
    ctx: Arc<GpuCtx>,
    /// CollectLeft: the build side is drained once and shared by every probe partition (`OnceAsync<JoinLeftData>`)
    shared_build: Arc<Mutex<Option<Arc<GpuJoin>>>>,
    /// probe partitions still running; the last one emits the unmatched build rows (CollectLeft + LeftSemi only, see rule.rs)
    metrics: ExecutionPlanMetricsSet,
}
impl fmt::Debug for GpuJoin {
    fn fmt(&self, f: &mut fmt::Formatter) -> fmt::Result {
        write!(f, "GpuJoin")
    }
}

impl GpuHashJoinExec {
    pub fn try_new(cpu: Arc<HashJoinExec>, ctx: Arc<GpuCtx>) -> Result<Self> {
        // CollectLeft keeps ONE visited bitmap for all probe partitions here, the reference one per partition
        // (hash_join.rs:1172-1190); the optimizer never plans CollectLeft for Left / Full / LeftAnti
        // (core/src/physical_optimizer/join_selection.rs:283-299) and neither does the rule.
        if *cpu.partition_mode() == PartitionMode::CollectLeft
            && matches!(cpu.join_type(), JoinType::Left | JoinType::Full | JoinType::LeftAnti | JoinType::LeftSemi)
            && cpu.right().output_partitioning().partition_count() > 1
        {
            return Err(DataFusionError::NotImplemented("CollectLeft with unmatched-build output over several probe partitions".into()));
        }
        drop(GpuJoin::create(&ctx, &cpu)?); // validates key / payload types at planning time
        Ok(Self { cpu, ctx, shared_build: Arc::new(Mutex::new(None)), metrics: ExecutionPlanMetricsSet::new() })
    }
    pub fn cpu(&self) -> &Arc<HashJoinExec> {
        &self.cpu
    }
}

impl DisplayAs for GpuHashJoinExec {
    fn fmt_as(&self, t: DisplayFormatType, f: &mut fmt::Formatter) -> fmt::Result {
        write!(f, "Gpu")?;
        self.cpu.fmt_as(t, f)
    }
}

impl ExecutionPlan for GpuHashJoinExec {
    fn as_any(&self) -> &dyn Any {
        self
    }
    fn properties(&self) -> &PlanProperties {
        self.cpu.properties() // build_join_schema (joins/utils.rs:657-729), output partitioning, maintains probe order
    }
    fn required_input_distribution(&self) -> Vec<Distribution> {
        self.cpu.required_input_distribution() // [SinglePartition, Unspecified] | [Hash(left), Hash(right)] (:514-533)
    }
    fn maintains_input_order(&self) -> Vec<bool> {
        self.cpu.maintains_input_order()
    }
    fn children(&self) -> Vec<Arc<dyn ExecutionPlan>> {
        self.cpu.children()
    }
    fn with_new_children(self: Arc<Self>, c: Vec<Arc<dyn ExecutionPlan>>) -> Result<Arc<dyn ExecutionPlan>> {
        let cpu = Arc::new(HashJoinExec::try_new(
            c[0].clone(),
            c[1].clone(),
            self.cpu.on().to_vec(),
            self.cpu.filter().cloned(),
            self.cpu.join_type(),
            *self.cpu.partition_mode(),
            self.cpu.null_equals_null(),
        )?);
        Ok(Arc::new(GpuHashJoinExec::try_new(cpu, self.ctx.clone())?))
    }
    fn execute(&self, partition: usize, tctx: Arc<TaskContext>) -> Result<SendableRecordBatchStream> {
        let collect_left = *self.cpu.partition_mode() == PartitionMode::CollectLeft;
        let left_parts = self.cpu.left().output_partitioning().partition_count();
        let right_parts = self.cpu.right().output_partitioning().partition_count();
        if !collect_left && left_parts != right_parts {
            return Err(DataFusionError::Internal(format!(
                "Invalid HashJoinExec, partition count mismatch {left_parts}!={right_parts}, consider using RepartitionExec"
            ))); // hash_join.rs:584-590
        }
        let mut build_in = if collect_left {
            if left_parts != 1 {
                return Err(DataFusionError::Plan("CollectLeft requires a single build partition".into()));
            }
            self.cpu.left().execute(0, tctx.clone())?
        } else {
            self.cpu.left().execute(partition, tctx.clone())?
        };
        let mut probe_in = self.cpu.right().execute(partition, tctx.clone())?;
        let (cpu, ctx, shared) = (self.cpu.clone(), self.ctx.clone(), self.shared_build.clone());
        let batch_size = tctx.session_config().batch_size();
        let schema: SchemaRef = self.schema();
        let out_schema = schema.clone();
        let join_time = MetricBuilder::new(&self.metrics).subset_time("join_time", partition);
        let output_rows = MetricBuilder::new(&self.metrics).output_rows(partition);
        let last_partition = partition + 1 == right_parts;
        let stream = futures::stream::once(async move {
            // collect_left_input: the first partition to get here builds, the others reuse (CollectLeft)
            let existing = if collect_left { shared.lock().clone() } else { None };
            let join = match existing {
                Some(j) => j,
                None => {
                    let j = GpuJoin::create(&ctx, &cpu)?;
                    while let Some(b) = build_in.next().await {
                        j.build_push(&b?)?;
                    }
                    j.build_finish()?;
                    let j = Arc::new(j);
                    if collect_left {
                        *shared.lock() = Some(j.clone());
                    }
                    j
                }
            };
            let mut out: Vec<RecordBatch> = vec![];
            while let Some(b) = probe_in.next().await {
                let r = join.probe(&b?)?; // every row the reference emits for this probe batch (its <= batch_size sub-batches)
                out.extend(slice_batches(RecordBatch::try_new(out_schema.clone(), r.columns().to_vec())?, batch_size));
            }
            // process_unmatched_build_batch: after the probe side is exhausted (Partitioned: per partition; CollectLeft with a
            // single probe partition, the only shape try_new accepts for these join types: here)
            if !collect_left || last_partition {
                if let Some(u) = join.emit_unmatched()? {
                    out.extend(slice_batches(RecordBatch::try_new(out_schema.clone(), u.columns().to_vec())?, batch_size));
                }
            }
            let m = join.metrics();
            join_time.add_duration(std::time::Duration::from_nanos(m.join_time_ns as u64));
            output_rows.add(out.iter().map(|b| b.num_rows()).sum());
            Ok::<_, DataFusionError>(futures::stream::iter(out.into_iter().map(Ok)))
        })
        .try_flatten();
        Ok(Box::pin(RecordBatchStreamAdapter::new(schema, stream)))
    }
    fn metrics(&self) -> Option<MetricsSet> {
        Some(self.metrics.clone_inner())
    }
}

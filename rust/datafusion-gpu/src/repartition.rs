//! `GpuRepartitionExec`: `RepartitionExec(Partitioning::Hash(exprs, n))` (physical-plan/src/repartition/mod.rs:133-217,
//! 442-580) whose `BatchPartitioner` is `dfgpu_partition_hash` (stable: rows keep their input order inside a partition, as
//! `take` does) and, when partition i lives on rank i of the context's communicator, whose channels are
//! `dfgpu_exchange_all_to_all` over NVLink.
//!
//! The row hash is the library's own ("dfhash v1", `dfgpu_hash_contract() == 1`), not ahash with the fixed seeds of
//! repartition/mod.rs:115: rows land in other partitions than the CPU operator would put them.  Only co-location is
//! observable, so a plan is correct as long as BOTH producers of a co-partitioned pair (the two inputs of a Partitioned
//! join; Partial -> Final) use the same contract; `GpuOffloadRule` replaces such pairs together or not at all.
use std::any::Any;
use std::fmt;
use std::ptr;
use std::sync::Arc;

use arrow::record_batch::RecordBatch;
use datafusion_common::{DataFusionError, Result};
use datafusion_execution::{SendableRecordBatchStream, TaskContext};
use datafusion_physical_expr::expressions::Column;
use datafusion_physical_expr::Partitioning;
use datafusion_physical_plan::metrics::{ExecutionPlanMetricsSet, MetricsSet};
use datafusion_physical_plan::repartition::RepartitionExec;
use datafusion_physical_plan::stream::RecordBatchStreamAdapter;
use datafusion_physical_plan::{DisplayAs, DisplayFormatType, ExecutionPlan, PlanProperties};
use futures::{StreamExt, TryStreamExt};
use parking_lot::Mutex;

use crate::ctx::{export_schema, slice_batches, DeviceBatch, GpuCtx};
use crate::sys::*;

#[derive(Debug)]
pub struct GpuRepartitionExec {
    cpu: Arc<RepartitionExec>,
    ctx: Arc<GpuCtx>,
    key_cols: Vec<i32>,
    n_parts: usize,
    /// output partitions, filled by whoever polls first (the reference keeps its channels behind a Mutex too, mod.rs:301)
    state: Arc<Mutex<Option<Vec<Vec<RecordBatch>>>>>,
    metrics: ExecutionPlanMetricsSet,
}

impl GpuRepartitionExec {
    pub fn try_new(cpu: Arc<RepartitionExec>, ctx: Arc<GpuCtx>) -> Result<Self> {
        let (exprs, n_parts) = match cpu.partitioning() {
            Partitioning::Hash(e, n) => (e.clone(), *n),
            other => return Err(DataFusionError::NotImplemented(format!("GPU repartitioning of {other:?}"))), // mod.rs:118
        };
        let key_cols = exprs
            .iter()
            .map(|e| {
                e.as_any()
                    .downcast_ref::<Column>()
                    .map(|c| c.index() as i32)
                    .ok_or_else(|| DataFusionError::NotImplemented("non-column hash expression on the GPU path".into()))
            })
            .collect::<Result<Vec<_>>>()?;
        if ctx.options.world > 1 && n_parts != ctx.options.world as usize {
            return Err(DataFusionError::Plan(format!("{n_parts} hash partitions over {} ranks", ctx.options.world)));
        }
        Ok(Self { cpu, ctx, key_cols, n_parts, state: Arc::new(Mutex::new(None)), metrics: ExecutionPlanMetricsSet::new() })
    }

    /// One input batch -> its n_parts sub-batches (None for empty ones: they are not produced, mod.rs:196-214).
    fn partition(&self, batch: &RecordBatch) -> Result<Vec<Option<DeviceBatch>>> {
        let d = DeviceBatch::from_record_batch(&self.ctx, batch)?;
        let mut parts = vec![ptr::null_mut(); self.n_parts];
        check(unsafe {
            dfgpu_partition_hash(self.ctx.ptr, d.ptr, self.key_cols.as_ptr(), self.key_cols.len() as i32, self.n_parts as i32, parts.as_mut_ptr())
        })?;
        Ok(parts.into_iter().map(|p| DeviceBatch::from_raw(&self.ctx, p)).collect())
    }
}

impl DisplayAs for GpuRepartitionExec {
    fn fmt_as(&self, t: DisplayFormatType, f: &mut fmt::Formatter) -> fmt::Result {
        write!(f, "Gpu")?;
        self.cpu.fmt_as(t, f)
    }
}

impl ExecutionPlan for GpuRepartitionExec {
    fn as_any(&self) -> &dyn Any {
        self
    }
    fn properties(&self) -> &PlanProperties {
        self.cpu.properties()
    }
    fn children(&self) -> Vec<Arc<dyn ExecutionPlan>> {
        self.cpu.children()
    }
    fn with_new_children(self: Arc<Self>, mut c: Vec<Arc<dyn ExecutionPlan>>) -> Result<Arc<dyn ExecutionPlan>> {
        let cpu = Arc::new(RepartitionExec::try_new(c.swap_remove(0), self.cpu.partitioning().clone())?);
        Ok(Arc::new(GpuRepartitionExec::try_new(cpu, self.ctx.clone())?))
    }
    fn execute(&self, partition: usize, tctx: Arc<TaskContext>) -> Result<SendableRecordBatchStream> {
        let world = self.ctx.options.world as usize;
        let input = self.cpu.input().clone();
        let n_in = input.output_partitioning().partition_count();
        let schema = self.schema();
        let batch_size = tctx.session_config().batch_size();
        // Sharded: this process IS output partition `rank`; one process, all partitions: pull_from_input for every input
        // partition on first poll (mod.rs:684-760), hand out partition `partition`.
        let this = GpuRepartitionExec {
            cpu: self.cpu.clone(),
            ctx: self.ctx.clone(),
            key_cols: self.key_cols.clone(),
            n_parts: self.n_parts,
            state: self.state.clone(),
            metrics: self.metrics.clone(),
        };
        let out_schema = schema.clone();
        let stream = futures::stream::once(async move {
            let ready = this.state.lock().is_some();
            if !ready {
                let mut outs: Vec<Vec<RecordBatch>> = vec![vec![]; this.n_parts];
                let ffi_schema = export_schema(&out_schema)?;
                for i in 0..n_in {
                    let mut s = input.execute(i, tctx.clone())?;
                    while let Some(b) = s.next().await {
                        let parts = this.partition(&b?)?;
                        if world > 1 {
                            // the channel hop: parts[r] goes to rank r; what every rank sent to this one comes back concatenated
                            let raw: Vec<*mut dfgpu_batch> = parts.iter().map(|p| p.as_ref().map_or(ptr::null_mut(), |d| d.ptr)).collect();
                            let mut recv = ptr::null_mut();
                            check(unsafe { dfgpu_exchange_all_to_all(this.ctx.ptr, raw.as_ptr(), &ffi_schema, &mut recv) })?;
                            if let Some(d) = DeviceBatch::from_raw(&this.ctx, recv) {
                                outs[this.ctx.options.rank as usize].push(d.to_record_batch()?);
                            }
                        } else {
                            for (p, d) in parts.into_iter().enumerate() {
                                if let Some(d) = d {
                                    outs[p].push(d.to_record_batch()?);
                                }
                            }
                        }
                    }
                }
                let mut g = this.state.lock();
                if g.is_none() {
                    *g = Some(outs);
                }
            }
            let mine = std::mem::take(&mut this.state.lock().as_mut().expect("filled above")[partition]);
            let mut out = vec![];
            for b in mine {
                out.extend(slice_batches(RecordBatch::try_new(out_schema.clone(), b.columns().to_vec())?, batch_size));
            }
            Ok::<_, DataFusionError>(futures::stream::iter(out.into_iter().map(Ok)))
        })
        .try_flatten();
        Ok(Box::pin(RecordBatchStreamAdapter::new(schema, stream)))
    }
    fn metrics(&self) -> Option<MetricsSet> {
        Some(self.metrics.clone_inner())
    }
}

//! `GpuOffloadRule`: the `PhysicalOptimizerRule` (core/src/physical_optimizer/optimizer.rs:39-55) that swaps the CPU hash
//! operators for their device counterparts.  Register it last:
//!
//! ```ignore
//! let ctx = GpuCtx::try_new(GpuOptions::default())?;
//! let state = SessionState::new_with_config_rt(config, runtime).add_physical_optimizer_rule(Arc::new(GpuOffloadRule::new(ctx)));
//! ```
//! (core/src/execution/context/mod.rs:1563-1568).  `schema_check()` is true: the replacements report the CPU nodes' own
//! schemas.  A node the device path does not take (`NotImplemented` from its constructor) stays as it is — there is no
//! fallback inside the library.
use std::sync::Arc;

use datafusion_common::config::ConfigOptions;
use datafusion_common::tree_node::{Transformed, TreeNode};
use datafusion_common::{DataFusionError, Result};
use datafusion::physical_optimizer::PhysicalOptimizerRule;
use datafusion_physical_expr::Partitioning;
use datafusion_physical_plan::aggregates::{AggregateExec, AggregateMode};
use datafusion_physical_plan::joins::{HashJoinExec, PartitionMode};
use datafusion_physical_plan::repartition::RepartitionExec;
use datafusion_physical_plan::ExecutionPlan;

use crate::aggregate::{func_code, GpuAggregateExec};
use crate::ctx::GpuCtx;
use crate::hash_join::GpuHashJoinExec;
use crate::repartition::GpuRepartitionExec;

pub struct GpuOffloadRule {
    ctx: Arc<GpuCtx>,
}

impl GpuOffloadRule {
    pub fn new(ctx: Arc<GpuCtx>) -> Self {
        // every partitioning decision of the replaced operators comes from ONE row hash; refuse a library that changed it
        // under a plan cache (INTEGRATION.md "hash contract")
        assert_eq!(GpuCtx::hash_contract(), 1, "libdfgpu row-hash contract changed");
        Self { ctx }
    }

    /// `NotImplemented` = "not on the device path": keep the CPU node.
    fn keep_on_cpu<T>(r: Result<T>) -> Result<Option<T>> {
        match r {
            Ok(v) => Ok(Some(v)),
            Err(DataFusionError::NotImplemented(_)) => Ok(None),
            Err(e) => Err(e),
        }
    }

    fn aggregate(&self, a: &AggregateExec) -> Result<Option<Arc<dyn ExecutionPlan>>> {
        if !a.aggr_expr().iter().all(|e| func_code(e).is_some()) {
            return Ok(None);
        }
        // Sharded plans: FinalPartitioned <- RepartitionExec(Hash(group keys), world) <- Partial becomes ONE operator that
        // exchanges partial-state records itself (dfgpu_agg_exchange; core/src/physical_planner.rs:814-849 is the pattern).
        if self.ctx.options.world > 1 && *a.mode() == AggregateMode::FinalPartitioned {
            if let Some(rep) = a.input().as_any().downcast_ref::<RepartitionExec>() {
                let hash_n = matches!(rep.partitioning(), Partitioning::Hash(_, n) if *n == self.ctx.options.world as usize);
                // (the Partial below may already have been replaced by this bottom-up pass)
                let partial = rep
                    .input()
                    .as_any()
                    .downcast_ref::<GpuAggregateExec>()
                    .map(|g| g.cpu().clone())
                    .or_else(|| rep.input().as_any().downcast_ref::<AggregateExec>().map(|p| Arc::new(clone_agg(p))));
                if let (true, Some(p)) = (hash_n, partial) {
                    if *p.mode() == AggregateMode::Partial {
                        let single = Arc::new(AggregateExec::try_new(
                            AggregateMode::SinglePartitioned, // output schema = the Final node's (result fields)
                            p.group_expr().clone(),
                            p.aggr_expr().to_vec(),
                            p.filter_expr().to_vec(),
                            p.input().clone(),
                            p.input_schema(),
                        )?);
                        if let Some(g) = Self::keep_on_cpu(GpuAggregateExec::try_new(single, self.ctx.clone(), true))? {
                            return Ok(Some(Arc::new(g)));
                        }
                    }
                }
            }
        }
        Ok(Self::keep_on_cpu(GpuAggregateExec::try_new(Arc::new(clone_agg(a)), self.ctx.clone(), false))?.map(|g| Arc::new(g) as _))
    }

    fn join(&self, j: &HashJoinExec) -> Result<Option<Arc<dyn ExecutionPlan>>> {
        // CollectLeft shares one visited bitmap between probe partitions here; the optimizer never plans it for the join
        // types that emit unmatched build rows (join_selection.rs:283-299), and GpuHashJoinExec::try_new refuses the rest.
        if *j.partition_mode() == PartitionMode::Auto {
            return Ok(None); // JoinSelection has not run yet
        }
        // Partitioned: both children are RepartitionExec(Hash) that must agree on the row hash: replace both or neither.
        let mut children = vec![j.left().clone(), j.right().clone()];
        if *j.partition_mode() == PartitionMode::Partitioned {
            let mut swapped = vec![];
            for c in &children {
                match c.as_any().downcast_ref::<RepartitionExec>() {
                    Some(r) => match Self::keep_on_cpu(GpuRepartitionExec::try_new(Arc::new(clone_rep(r)?), self.ctx.clone()))? {
                        Some(g) => swapped.push(Arc::new(g) as Arc<dyn ExecutionPlan>),
                        None => return Ok(None),
                    },
                    // already co-partitioned by something else (e.g. a bucketed scan): its hash is the CPU's, keep both on the
                    // CPU contract unless it already is one of ours
                    None if c.as_any().is::<GpuRepartitionExec>() => swapped.push(c.clone()),
                    None => return Ok(None),
                }
            }
            children = swapped;
        }
        let cpu = Arc::new(HashJoinExec::try_new(
            children[0].clone(),
            children[1].clone(),
            j.on().to_vec(),
            j.filter().cloned(), // residual JoinFilter: dfgpu_join_probe_candidates / _commit, never a post-filter
            j.join_type(),
            *j.partition_mode(),
            j.null_equals_null(),
        )?);
        Ok(Self::keep_on_cpu(GpuHashJoinExec::try_new(cpu, self.ctx.clone()))?.map(|g| Arc::new(g) as _))
    }
}

fn clone_agg(a: &AggregateExec) -> AggregateExec {
    AggregateExec::try_new(*a.mode(), a.group_expr().clone(), a.aggr_expr().to_vec(), a.filter_expr().to_vec(), a.input().clone(), a.input_schema())
        .expect("re-creating an AggregateExec from its own parts")
}
fn clone_rep(r: &RepartitionExec) -> Result<RepartitionExec> {
    RepartitionExec::try_new(r.input().clone(), r.partitioning().clone())
}

impl PhysicalOptimizerRule for GpuOffloadRule {
    fn name(&self) -> &str {
        "gpu_offload"
    }
    fn schema_check(&self) -> bool {
        true
    }
    fn optimize(&self, plan: Arc<dyn ExecutionPlan>, _config: &ConfigOptions) -> Result<Arc<dyn ExecutionPlan>> {
        let out = plan.transform_up(&|p: Arc<dyn ExecutionPlan>| {
            let replaced = if let Some(a) = p.as_any().downcast_ref::<AggregateExec>() {
                self.aggregate(a)?
            } else if let Some(j) = p.as_any().downcast_ref::<HashJoinExec>() {
                self.join(j)?
            } else {
                None
            };
            Ok(match replaced {
                Some(n) => Transformed::yes(n),
                None => Transformed::no(p),
            })
        })?;
        Ok(out.data)
    }
}

//! B200-native hash aggregation / hash join / hash repartition behind DataFusion's own operator interfaces.
//!
//! * [`sys`] — `extern "C"` declarations of `include/dfgpu.h`, one to one.
//! * [`ctx`] — safe handles (`GpuCtx`, `DeviceBatch`) and the Arrow C Data Interface hops.
//! * [`aggregate::GpuAggregateExec`] — `AggregateExec` / `GroupedHashAggregateStream` (physical-plan/src/aggregates/).
//! * [`hash_join::GpuHashJoinExec`] — `HashJoinExec` (physical-plan/src/joins/hash_join.rs), all eight join types, residual
//!   `JoinFilter` through `dfgpu_join_probe_candidates` / `_commit`.
//! * [`repartition::GpuRepartitionExec`] — `RepartitionExec(Partitioning::Hash)` (physical-plan/src/repartition/mod.rs).
//! * [`rule::GpuOffloadRule`] — the `PhysicalOptimizerRule` that swaps them in
//!   (`SessionState::add_physical_optimizer_rule`, core/src/execution/context/mod.rs:1563-1568).
//!
//! `PhysicalExpr::evaluate` (group keys, aggregate arguments, FILTER clauses, join keys, join filters) stays on the CPU;
//! the operators hand evaluated arrays to the device and take `RecordBatch`es back.  There is no CPU fallback inside the
//! library: what it does not implement is left as the CPU operator by the rule.
pub mod aggregate;
pub mod ctx;
pub mod hash_join;
pub mod repartition;
pub mod rule;
pub mod sys;

pub use aggregate::GpuAggregateExec;
pub use ctx::{DeviceBatch, GpuCtx, GpuOptions};
pub use hash_join::GpuHashJoinExec;
pub use repartition::GpuRepartitionExec;
pub use rule::GpuOffloadRule;

"""B200-native hash aggregation / hash join / hash repartition for DataFusion-shaped plans.

The product is ``libdfgpu.so`` (hand-written CUDA for sm_100a behind the C ABI in ``include/dfgpu.h``);
this package is the thin host-side mirror of the reference operator interface used by tests and bench.
"""
from ._native import DfGpuError, LIB_PATH, SYMBOLS  # noqa: F401
from .exec import (AggFunc, AggregateExpr, AggregateMode, DeviceBatch, GpuAggregateExec, GpuContext,  # noqa: F401
                   GpuCoalesceBatchesExec, GpuHashJoinExec, GpuRepartitionExec, JoinType, PartitionMode, all_gather, create_hashes, slice_batches)

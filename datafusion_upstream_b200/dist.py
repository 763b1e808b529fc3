"""Host-side plumbing for the sharded (N > 1) path: one process per GPU, `torch.distributed` only for
bootstrap and barriers, the data path is NCCL inside libdfgpu (`dfgpu_exchange_all_to_all`).

Sharding mirrors the reference's own plan shape: input rows are range-sharded across ranks,
`AggregateExec(Partial)` runs per rank, `RepartitionExec(Hash(group keys), N)` becomes the all-to-all, and
`AggregateExec(FinalPartitioned)` runs on the rows a rank owns (`core/src/physical_planner.rs:814-849`).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import pyarrow as pa


def shard_range(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Rows [lo, hi) of rank `rank`: contiguous, covering, sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("rank/world out of range")
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def owner_of(hashes, world: int):
    """Owner rank of each row = hash % world, the rule of `BatchPartitioner` (repartition/mod.rs:184-187)."""
    return hashes % world


def bootstrap_nccl_id(dist, make_id) -> Optional[bytes]:
    """Rank 0 creates the ncclUniqueId (`make_id()`), everyone receives it over the bootstrap group."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return None
    box = [make_id() if dist.get_rank() == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def final_specs(aggr_exprs: Sequence) -> List:
    """Final-mode aggregate list for a Partial-mode list: same functions and names, arguments resolved by
    position in the Partial output (aggregates/mod.rs:1038-1067)."""
    from .exec import AggregateExpr
    return [AggregateExpr(e.func, None, e.name) for e in aggr_exprs]


def sharded_aggregate(ctx, schema: pa.Schema, group_by: Sequence, aggr_exprs: Sequence, local_batches, world: int, to_host: bool = True):
    """Partial on the local shard -> hash exchange -> FinalPartitioned on the owned keys.  With world == 1 the
    exchange degenerates to a concat, so the same code path is exercised on one GPU."""
    from .exec import AggregateMode, GpuAggregateExec, GpuRepartitionExec
    part = GpuAggregateExec(ctx, AggregateMode.Partial, schema, group_by, aggr_exprs)
    for b in local_batches:
        part.push(b)
    st = part.emit_device()
    nkeys = len(group_by)
    rp = GpuRepartitionExec(ctx, part.schema, list(range(nkeys)), world)
    recv = rp.exchange(rp.partition_device(st))
    fin = GpuAggregateExec(ctx, AggregateMode.FinalPartitioned, part.schema, list(range(nkeys)), final_specs(aggr_exprs))
    if recv is not None:
        fin.push(recv)
    return fin.emit() if to_host else fin.emit_device()


def fused_sharded_aggregate(ctx, schema: pa.Schema, group_by: Sequence, aggr_exprs: Sequence, local_batches, world: int, to_host: bool = True):
    """The same plan fused into one operator per rank (`dfgpu_agg_exchange`): level-1 pre-aggregation of the local
    shard, partial-state records partitioned by key hash and exchanged over NCCL (one contiguous transfer per peer),
    final merge of the owned partitions.  Falls back to the three-operator plan when the operator cannot exchange at
    the record level (the decision depends only on schema and options, so every rank takes the same branch)."""
    from ._native import DfGpuError
    from .exec import AggregateMode, GpuAggregateExec
    op = GpuAggregateExec(ctx, AggregateMode.Single, schema, group_by, aggr_exprs)
    local_batches = list(local_batches)
    for b in local_batches:
        op.push(b)
    try:
        op.exchange()
    except DfGpuError as e:
        if e.kind != "NotImplemented":
            raise
        op.close()
        return sharded_aggregate(ctx, schema, group_by, aggr_exprs, local_batches, world, to_host)
    return op.emit() if to_host else op.emit_device()

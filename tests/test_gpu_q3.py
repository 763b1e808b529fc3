"""TPC-H Q3-shaped operator chain (BASELINE config C3) on synthetic tables, end to end through the
C ABI and against the oracle:

    customer(filtered) JOIN orders(filtered) ON c_custkey = o_custkey            HashJoinExec Inner / CollectLeft
      -> project (o_orderkey, o_orderdate, o_shippriority)
      JOIN lineitem(filtered) ON o_orderkey = l_orderkey                          HashJoinExec Inner / CollectLeft
      -> project (l_orderkey, o_orderdate, o_shippriority, l_extendedprice, l_discount)
      -> AggregateExec(Partial) per input partition -> hash repartition -> AggregateExec(FinalPartitioned)
         GROUP BY l_orderkey, o_orderdate, o_shippriority  SUM(l_extendedprice), SUM(l_discount), COUNT(*)

The reference plan is sqllogictest/test_files/tpch/q3.slt.part:72-92; the revenue expression itself is a
PhysicalExpr outside the path, so the chain sums the Decimal128 payload columns as they are.  Besides the oracle
the result is cross-checked with an independent numpy group-by.
"""
import numpy as np
import pyarrow as pa
import pytest

import oracle
from gpu_helpers import gpu_agg, gpu_ctx, oracle_agg
from util import assert_rows_equal, rows_of

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = gpu_ctx()
    yield c
    c.close()


def make_q3_tables(rng, n_cust, n_orders, lines_max=7):
    """Post-filter inputs of the three scans, shaped like the generator of SURVEY 8d."""
    custkey = np.arange(1, n_cust + 1, dtype=np.int64)
    customer = pa.record_batch([pa.array(custkey[rng.random(n_cust) < 0.2])], names=["c_custkey"])
    ok = np.arange(n_orders, dtype=np.int64)
    o_orderkey = (ok // 8) * 32 + ok % 8 + 1                      # sparse order keys: 8 of every 32
    valid_cust = custkey[custkey % 3 != 0]
    o_custkey = valid_cust[rng.integers(0, len(valid_cust), n_orders)]
    o_date = rng.integers(8035, 10440, n_orders).astype(np.int32)  # 1992-01-01 .. 1998-08-02
    keep = o_date < 9204                                           # o_orderdate < 1995-03-15 keeps ~48 %
    orders = pa.record_batch([pa.array(o_orderkey[keep]), pa.array(o_custkey[keep]), pa.array(o_date[keep], pa.int32()).cast(pa.date32()),
                              pa.array(np.zeros(int(keep.sum()), np.int32))],
                             names=["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"])
    nlines = rng.integers(1, lines_max + 1, n_orders)
    l_orderkey = np.repeat(o_orderkey, nlines)
    l_ship = np.repeat(o_date, nlines) + rng.integers(1, 122, len(l_orderkey))
    lkeep = l_ship > 9204                                          # l_shipdate > 1995-03-15 keeps ~54 %
    n_l = int(lkeep.sum())
    price = pa.array([int(v) for v in rng.integers(90000, 10495001, n_l)], pa.decimal128(15, 2))
    disc = pa.array([int(v) for v in rng.integers(0, 11, n_l)], pa.decimal128(15, 2))
    lineitem = pa.record_batch([pa.array(l_orderkey[lkeep]), price, disc], names=["l_orderkey", "l_extendedprice", "l_discount"])
    return customer, orders, lineitem


def _dec_array(values, typ):
    import decimal
    return pa.array([decimal.Decimal(int(v)).scaleb(-typ.scale) for v in values], typ)


SPECS = [("sum", "l_extendedprice", "SUM(l_extendedprice)", None), ("sum", "l_discount", "SUM(l_discount)", None),
         ("count", None, "COUNT(*)", None)]
GROUP = ["l_orderkey", "o_orderdate", "o_shippriority"]


def gpu_chain(ctx, customer, orders, lineitem, probe_chunks=3, partitions=3):
    import datafusion_upstream_b200 as df
    j1 = df.GpuHashJoinExec(ctx, customer.schema, orders.schema, [("c_custkey", "o_custkey")], df.JoinType.Inner, df.PartitionMode.CollectLeft)
    j1.build([customer])
    o1 = j1.probe_device(orders).project(["o_orderkey", "o_orderdate", "o_shippriority"])
    j2 = df.GpuHashJoinExec(ctx, o1.schema, lineitem.schema, [("o_orderkey", "l_orderkey")], df.JoinType.Inner, df.PartitionMode.CollectLeft)
    j2.build([o1])
    step = -(-lineitem.num_rows // probe_chunks)
    joined = [j2.probe_device(lineitem.slice(i, step)).project(GROUP + ["l_extendedprice", "l_discount"])
              for i in range(0, lineitem.num_rows, step)]
    # one Partial aggregation per probe batch "partition", then the hash exchange, then FinalPartitioned per partition
    exprs = [df.AggregateExpr(df.AggFunc.Sum, "l_extendedprice", SPECS[0][2]), df.AggregateExpr(df.AggFunc.Sum, "l_discount", SPECS[1][2]),
             df.AggregateExpr(df.AggFunc.Count, None, SPECS[2][2])]
    states = []
    for jb in joined:
        part = df.GpuAggregateExec(ctx, df.AggregateMode.Partial, jb.schema, GROUP, exprs)
        part.push(jb)
        states.append(part.emit_device())
    sch = states[0].schema
    rp = df.GpuRepartitionExec(ctx, sch, [0, 1, 2], partitions)
    by_part = [[] for _ in range(partitions)]
    for st in states:
        for p, b in enumerate(rp.partition_device(st)):
            if b is not None:
                by_part[p].append(b)
    outs = []
    for p in range(partitions):
        fin = df.GpuAggregateExec(ctx, df.AggregateMode.FinalPartitioned, sch, [0, 1, 2], [df.AggregateExpr(e.func, None, e.name) for e in exprs])
        for b in by_part[p]:
            fin.push(b)
        outs.append(fin.emit())
    return pa.Table.from_batches(outs), sum(j.num_rows for j in joined), o1.num_rows


def oracle_chain(customer, orders, lineitem):
    t1 = oracle.hash_join("inner", [customer], [orders], [(0, 1)])
    o1 = t1.select(["o_orderkey", "o_orderdate", "o_shippriority"]).combine_chunks().to_batches()
    t2 = oracle.hash_join("inner", o1, [lineitem], [(0, 0)], left_schema=t1.select(["o_orderkey", "o_orderdate", "o_shippriority"]).schema)
    joined = t2.select(GROUP + ["l_extendedprice", "l_discount"]).combine_chunks().to_batches()
    return oracle_agg("single", joined, GROUP, SPECS, schema=t2.select(GROUP + ["l_extendedprice", "l_discount"]).schema), t2.num_rows


@pytest.mark.parametrize("scale", [(3_000, 30_000), (20_000, 200_000)], ids=["tiny", "small"])
def test_q3_chain_vs_oracle(ctx, scale):
    rng = np.random.default_rng(scale[0])
    customer, orders, lineitem = make_q3_tables(rng, *scale)
    got, joined_rows, j1_rows = gpu_chain(ctx, customer, orders, lineitem)
    want, want_joined = oracle_chain(customer, orders, lineitem)
    assert joined_rows == want_joined
    assert_rows_equal(rows_of(got), rows_of(want), key_cols=[0, 1, 2], what="q3 chain")
    # independent cross-check: numpy group-by on the join keys
    ckeys = set(customer.column(0).to_numpy().tolist())
    omask = np.array([c in ckeys for c in orders.column(1).to_numpy()])
    assert j1_rows == int(omask.sum())
    okeys = orders.column(0).to_numpy()[omask]
    lk = lineitem.column(0).to_numpy()
    lmask = np.isin(lk, okeys)
    uniq, counts = np.unique(lk[lmask], return_counts=True)
    got_counts = dict(zip(got.column(0).to_pylist(), got.column(5).to_pylist()))
    assert got.num_rows == len(uniq)
    assert got_counts == dict(zip(uniq.tolist(), counts.tolist()))
    price = np.array([int(v.as_py().scaleb(2)) for v in lineitem.column(1)], dtype=object)[lmask]
    assert sum(int(v.as_py().scaleb(2)) for v in got.column(3)) == int(price.sum())


def test_q3_chain_partition_count_invariant(ctx):
    """The result does not depend on how the probe side is batched nor on the number of exchange partitions."""
    rng = np.random.default_rng(11)
    customer, orders, lineitem = make_q3_tables(rng, 2_000, 15_000)
    a, _, _ = gpu_chain(ctx, customer, orders, lineitem, probe_chunks=1, partitions=1)
    b, _, _ = gpu_chain(ctx, customer, orders, lineitem, probe_chunks=5, partitions=8)
    assert_rows_equal(rows_of(a), rows_of(b), key_cols=[0, 1, 2], what="q3 partitioning invariance")

"""Pins the CPU oracle to the reference's own literal test vectors (tests/golden/*.json, generated
from the reference sources by tests/golden/make_golden.py).  CPU-only."""
import decimal
import json
import os

import numpy as np
import pyarrow as pa
import pytest

import oracle
from oracle import AggSpec, OracleAgg
from util import assert_rows_equal, golden_filter, rows_of

G = os.path.join(os.path.dirname(__file__), "golden")
JOINS = json.load(open(os.path.join(G, "hash_join.json")))
AGGS = json.load(open(os.path.join(G, "aggregates.json")))
JT = {"Inner": "inner", "Left": "left", "Right": "right", "Full": "full", "LeftSemi": "left_semi",
      "RightSemi": "right_semi", "LeftAnti": "left_anti", "RightAnti": "right_anti"}


def batches(side):
    out = []
    for b in side["batches"]:
        out.append(pa.record_batch([pa.array(c, pa.int32()) for c in b["columns"]], names=b["names"]))
    return out


# the reference runs every join test with these batch sizes (hash_join.rs:1434)
@pytest.mark.parametrize("batch_size", [8192, 10, 5, 2, 1])
@pytest.mark.parametrize("case", JOINS, ids=[c["name"] for c in JOINS])
def test_join_golden(case, batch_size):
    left, right = batches(case["left"]), batches(case["right"])
    on = [(left[0].schema.names.index(l), right[0].schema.names.index(r)) for l, r in case["on"]]
    t = oracle.hash_join(JT[case["join_type"]], left, right, on, batch_size=batch_size,
                         force_collisions=case.get("force_collisions", False), filter=golden_filter(case))
    assert t.schema.names == case["expected_header"]
    want = [tuple(r) for r in case["expected"]]
    assert_rows_equal(rows_of(t), want, ordered=case["ordered"], what=case["name"])


def test_join_inner_two_batch_count():
    # hash_join.rs:1752-1758: ceil(3 / batch_size) output batches, +1 empty batch when batch_size == 1
    case = next(c for c in JOINS if c["name"] == "join_inner_two")
    left, right = batches(case["left"]), batches(case["right"])
    for bs in (8192, 10, 5, 2, 1):
        oracle.hash_join("inner", left, right, [(0, 0), (1, 1)], batch_size=bs)
        expect = -(-3 // bs) + (1 if bs == 1 else 0)
        assert oracle.hash_join.last_out_batches == expect


def some_data():
    return [pa.record_batch([pa.array(b["a"], pa.uint32()), pa.array(b["b"], pa.float64())], names=["a", "b"])
            for b in AGGS["some_data"]["batches"]]


def test_check_aggregates():
    g = AGGS["check_aggregates"]
    aggs = [AggSpec("avg", "AVG(b)", pa.float64())]
    p = OracleAgg("partial", ["a"], [pa.uint32()], aggs)
    for b in some_data():
        p.push([b.column(0)], [b.column(1)])
    st = p.emit()
    assert st.schema.names == g["partial_no_spill"]["header"]
    want = [(int(a), int(c), float(s)) for a, c, s in g["partial_no_spill"]["rows"]]
    assert_rows_equal(rows_of(st), want)
    f = OracleAgg("final", ["a"], [pa.uint32()], aggs)
    f.push([st.column(0)], [st.column(1), st.column(2)])
    out = f.emit()
    assert out.schema.names == g["final"]["header"]
    want = [(int(a), float(v)) for a, v in g["final"]["rows"]]
    assert_rows_equal(rows_of(out), want, rel=0.0)   # bit-exact: 2.3333333333333335, 3.6666666666666665


def test_check_grouping_sets():
    # one intern+update pass per grouping set with NULL-substituted key columns (aggregates/mod.rs:1183-1199)
    g = AGGS["check_grouping_sets"]
    aggs = [AggSpec("count", "COUNT(1)", None)]
    p = OracleAgg("partial", ["a", "b"], [pa.uint32(), pa.float64()], aggs)
    for b in some_data():
        for null_a, null_b in g["groups"]:
            a = pa.nulls(b.num_rows, pa.uint32()) if null_a else b.column(0)
            bb = pa.nulls(b.num_rows, pa.float64()) if null_b else b.column(1)
            p.push([a, bb], [None])
    st = p.emit()
    conv = lambda r: (None if r[0] is None else int(r[0]), None if r[1] is None else float(r[1]), int(r[2]))
    assert_rows_equal(rows_of(st), [conv(r) for r in g["partial_no_spill"]["rows"]])
    f = OracleAgg("final", ["a", "b"], [pa.uint32(), pa.float64()], aggs)
    f.push([st.column(0), st.column(1)], [st.column(2)])
    assert_rows_equal(rows_of(f.emit()), [conv(r) for r in g["final"]["rows"]])


@pytest.mark.parametrize("name", ["avg_decimal", "avg_decimal_with_nulls"])
def test_avg_decimal(name):
    g = AGGS[name]
    t = pa.decimal128(*g["input_type"])
    arr = pa.array([None if v is None else decimal.Decimal(v) for v in g["input"]], t)
    a = OracleAgg("single", ["k"], [pa.int8()], [AggSpec("avg", "avg", t)])
    a.push([pa.array([0] * len(arr), pa.int8())], [arr])
    out = a.emit()
    assert out.schema.field(1).type == pa.decimal128(*g["expected_type"])
    got = out.column(1)[0].as_py()
    assert got == decimal.Decimal(g["expected_unscaled"]).scaleb(-g["expected_type"][1])


def test_avg_decimal_all_nulls():
    t = pa.decimal128(10, 0)
    a = OracleAgg("single", ["k"], [pa.int8()], [AggSpec("avg", "avg", t)])
    a.push([pa.array([0] * 6, pa.int8())], [pa.nulls(6, t)])
    assert a.emit().column(1).to_pylist() == [None]     # average.rs:613-629


def test_decimal_slt():
    d = AGGS["decimal_slt"]
    t = pa.decimal128(*d["c1_type"])
    c1 = pa.array([decimal.Decimal(v) for v in d["c1"]], t)
    c4 = pa.array(d["c4"], pa.bool_())
    one = pa.array([0] * len(c1), pa.int8())
    a = OracleAgg("single", ["k"], [pa.int8()], [AggSpec("sum", "s", t), AggSpec("avg", "a", t)])
    a.push([one], [c1, c1])
    out = a.emit()
    # decimal.slt:112-124 — Decimal128(20,6) 0.00055 ; Decimal128(14,10) 0.0000366666 (truncation)
    assert out.schema.field(1).type == pa.decimal128(20, 6)
    assert out.schema.field(2).type == pa.decimal128(14, 10)
    assert out.column(1)[0].as_py() == decimal.Decimal(d["sum_c1"][0][-1])
    assert out.column(2)[0].as_py() == decimal.Decimal(d["avg_c1"][0][-1])
    a = OracleAgg("single", ["c1"], [t], [AggSpec("count", "c", None)])
    a.push([c1], [None])
    want = [(decimal.Decimal(v), int(c)) for c, v in d["count_by_c1"]]
    assert_rows_equal(rows_of(a.emit()), want)
    a = OracleAgg("single", ["c1", "c4"], [t, pa.bool_()], [AggSpec("count", "c", None)])
    a.push([c1, c4], [None])
    want = [(decimal.Decimal(v), b == "true", int(c)) for c, v, b in d["count_by_c1_c4"]]
    assert_rows_equal(rows_of(a.emit()), want)


def test_hash_properties():
    # the reference asserts only properties of create_hashes (hash_utils.rs:450-728)
    s = pa.array(["foo", None, "bar", "foo", "", None])
    h = oracle.create_hashes([s])
    assert h[0] == h[3] and h[1] == 0 and h[5] == 0 and h[0] != h[2] and h[4] != 0
    hd = oracle.create_hashes([s.dictionary_encode()])
    assert (h == hd).all()                                   # dict hashes == plain hashes (:520-562)
    i = pa.array([1, 2, None, 1], pa.int64())
    h2 = oracle.create_hashes([s.slice(0, 4), i])
    h2d = oracle.create_hashes([s.slice(0, 4).dictionary_encode(), i])
    assert (h2 == h2d).all() and h2[0] == h2[3] and h2[0] != h2[1]
    f = pa.array([0.0, -0.0, float("nan"), float("nan")])
    hf = oracle.create_hashes([f])
    assert hf[0] != hf[1] and hf[2] == hf[3]                 # bit-pattern hashing (:78-87)
    assert (oracle.create_hashes([i], force_collisions=True) == 0).all()


def test_partition_conserves_rows():
    rng = np.random.default_rng(0)
    b = pa.record_batch([pa.array(rng.integers(0, 50, 1000)), pa.array(rng.random(1000))], names=["k", "v"])
    parts = oracle.partition(b, [0], 7)
    assert sum(p.num_rows for p in parts if p is not None) == 1000   # repartition/mod.rs:1010-1031
    seen = {}
    for pi, p in enumerate(parts):
        if p is None:
            continue
        for k in p.column(0).to_pylist():
            assert seen.setdefault(k, pi) == pi


def test_emit_first_n_and_float_minmax_quirks():
    a = OracleAgg("single", ["k"], [pa.int64()], [AggSpec("max", "mx", pa.float64()), AggSpec("min", "mn", pa.float64()),
                                                   AggSpec("count", "c", pa.float64())])
    k = pa.array([1, 1, 2, 3, 3, 4], pa.int64())
    v = pa.array([float("nan"), float("nan"), float("-inf"), 1.5, None, None], pa.float64())
    a.push([k], [v, v, v])
    first = a.emit(2)
    rest = a.emit()
    import sys
    # min_max.rs:102-139: NaN never replaces the start value; start values are finite f64::MIN/MAX
    assert rows_of(first) == [(1, -sys.float_info.max, sys.float_info.max, 2), (2, -sys.float_info.max, float("-inf"), 1)]
    assert rows_of(rest) == [(3, 1.5, 1.5, 1), (4, None, None, 0)]


def _bit_slt_inputs():
    g = AGGS["bit_aggregate_slt"]
    cols = list(zip(*g["rows"]))
    c = {n: pa.array(list(v), pa.int16()) for n, v in zip(g["columns"][:3], cols[:3])}   # SMALLINT
    tag = pa.array(list(cols[3]), pa.string())
    specs = [(f, n) for f in ("bit_and", "bit_or", "bit_xor") for n in ("c1", "c2", "c3")]
    want = [tuple([r[-1]] + r[:-1]) for r in g["expected"]]
    return c, tag, specs, want


def test_bit_aggregate_slt():
    # aggregate.slt:1795-1813 — grouped BIT_AND / BIT_OR / BIT_XOR, NULL for a group without values
    c, tag, specs, want = _bit_slt_inputs()
    a = OracleAgg("single", ["tag"], [pa.string()], [AggSpec(f, f"{f}({n})", pa.int16()) for f, n in specs])
    a.push([tag], [c[n] for _, n in specs])
    out = a.emit()
    assert [f.type for f in out.schema][1:] == [pa.int16()] * 9
    assert_rows_equal(rows_of(out), want)
    # Partial (state == value, names `x[bit_and]`, bit_and_or_xor.rs:96-102) -> Final
    p = OracleAgg("partial", ["tag"], [pa.string()], [AggSpec(f, f"{f}({n})", pa.int16()) for f, n in specs])
    p.push([tag.slice(0, 2)], [c[n].slice(0, 2) for _, n in specs])
    s1 = p.emit()
    p.push([tag.slice(2)], [c[n].slice(2) for _, n in specs])
    s2 = p.emit()
    assert s1.schema.names[1] == "bit_and(c1)[bit_and]" and s1.schema.names[9] == "bit_xor(c3)[bit_xor]"
    f = OracleAgg("final", ["tag"], [pa.string()], [AggSpec(fn, f"{fn}({n})", pa.int16()) for fn, n in specs])
    for s in (s1, s2):
        f.push([s.column(0)], [s.column(i) for i in range(1, 10)])
    assert_rows_equal(rows_of(f.emit()), want)


@pytest.mark.parametrize("case", AGGS["bit_unit_tests"], ids=[c["name"] for c in AGGS["bit_unit_tests"]])
def test_bit_unit_vectors(case):
    # bit_and_or_xor.rs:708-787 (generic_test_op! known answers)
    t = pa.uint32() if case["unsigned"] else pa.int32()
    v = pa.array(case["input"], t)
    a = OracleAgg("single", ["k"], [pa.int8()], [AggSpec(case["op"], "r", t)])
    a.push([pa.array([0] * len(v), pa.int8())], [v])
    out = a.emit()
    assert out.schema.field(1).type == t
    assert out.column(1).to_pylist() == [case["expected"]]


def test_bool_and_or_groups_accumulator():
    # BooleanGroupsAccumulator (bool_op.rs:76-103): a new group's value bit starts at false for BOTH functions,
    # so grouped BOOL_AND is false whenever the group saw a value; NULL when it saw none.
    k = pa.array([1, 1, 2, 2, 3, 4, 4], pa.int32())
    v = pa.array([True, True, False, True, None, False, False], pa.bool_())
    a = OracleAgg("single", ["k"], [pa.int32()], [AggSpec("bool_and", "a", pa.bool_()), AggSpec("bool_or", "o", pa.bool_())])
    a.push([k], [v, v])
    out = a.emit()
    assert out.schema.field(1).type == pa.bool_()
    assert_rows_equal(rows_of(out), [(1, False, True), (2, False, True), (3, None, None), (4, False, False)])


# ---- Decimal128 arithmetic of ProjectionExec, pinned by decimal.slt:205-362 ----
DEC_ARITH = {
    "c1+1": [("col", 0), ("lit", 1, 20, 0), ("add", 0, 1)],        # the planner casts the Int64 literal to Decimal128(20, 0)
    "c1+c5": [("col", 0), ("col", 1), ("add", 0, 1)],
    "c1-1": [("col", 0), ("lit", 1, 20, 0), ("sub", 0, 1)],
    "c1-c5": [("col", 0), ("col", 1), ("sub", 0, 1)],
    "c1*20": [("col", 0), ("lit", 20, 20, 0), ("mul", 0, 1)],
    "c1*c5": [("col", 0), ("col", 1), ("mul", 0, 1)],
}


def decimal_simple():
    d = AGGS["decimal_slt"]
    c1 = pa.array([decimal.Decimal(v) for v in d["c1"]], pa.decimal128(*d["c1_type"]))
    c5 = pa.array([decimal.Decimal(v) for v in d["c5"]], pa.decimal128(*d["c5_type"]))
    return pa.record_batch([c1, c5], names=["c1", "c5"])


@pytest.mark.parametrize("name", list(DEC_ARITH))
def test_decimal_arithmetic_slt(name):
    g = AGGS["decimal_slt"]["arithmetic"][name]
    out = oracle.eval_decimal(decimal_simple(), DEC_ARITH[name])
    assert out.type == pa.decimal128(*g["type"])
    assert sorted(out.to_pylist()) == sorted(decimal.Decimal(v) for v in g["values_rowsort"])


def test_decimal_arithmetic_q1_types_nulls_overflow():
    # Q1's projections: price * (1 - discount) -> Decimal128(38, 4); that * (1 + tax) -> Decimal128(38, 6)  
    d152 = pa.decimal128(15, 2)
    b = pa.record_batch([pa.array([decimal.Decimal("100.00"), None, decimal.Decimal("-7.25")], d152),
                         pa.array([decimal.Decimal("0.05"), decimal.Decimal("0.01"), None], d152),
                         pa.array([decimal.Decimal("0.08")] * 3, d152)], names=["price", "disc", "tax"])
    disc_price = [("col", 0), ("lit", 1, 20, 0), ("col", 1), ("sub", 1, 2), ("mul", 0, 3)]
    out = oracle.eval_decimal(b, disc_price)
    assert out.type == pa.decimal128(38, 4) and out.to_pylist() == [decimal.Decimal("95.0000"), None, None]
    charge = disc_price + [("lit", 1, 20, 0), ("col", 2), ("add", 5, 6), ("mul", 4, 7)]
    out = oracle.eval_decimal(b, charge)
    assert out.type == pa.decimal128(38, 6) and out.to_pylist() == [decimal.Decimal("102.600000"), None, None]
    big = pa.record_batch([pa.array([decimal.Decimal(10) ** 37], pa.decimal128(38, 0))], names=["x"])
    with pytest.raises(oracle.OracleError, match="Overflow"):
        oracle.eval_decimal(big, [("col", 0), ("col", 0), ("mul", 0, 1)])
    with pytest.raises(oracle.OracleError, match="max scale"):
        oracle.eval_decimal(pa.record_batch([pa.array([decimal.Decimal("0.1")], pa.decimal128(38, 20))], names=["x"]),
                            [("col", 0), ("col", 0), ("mul", 0, 1)])


# ---- sum.rs / average.rs / count.rs / min_max.rs unit-test known answers (values, NULL handling, result types) ----
PA_T = {"int32": pa.int32(), "uint32": pa.uint32(), "int64": pa.int64(), "uint64": pa.uint64(), "float32": pa.float32(),
        "float64": pa.float64()}


@pytest.mark.parametrize("case", AGGS["simple_unit_tests"], ids=[c["name"] for c in AGGS["simple_unit_tests"]])
def test_simple_unit_vectors(case):
    t = PA_T[case["input_type"]]
    v = pa.array(case["input"], t)
    a = OracleAgg("single", ["k"], [pa.int8()], [AggSpec(case["func"], "r", t)])
    a.push([pa.array([0] * len(v), pa.int8())], [v])
    out = a.emit()
    assert out.schema.field(1).type == PA_T[case["expected_type"]], case["source"]
    assert out.column(1).to_pylist() == [case["expected"]], case["source"]

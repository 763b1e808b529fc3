"""-m gpu parity tests for the hash join: CUDA path vs the reference's golden tables
(tests/golden/hash_join.json, from hash_join.rs:1553-2943) and vs the CPU oracle (differential,
like core/tests/fuzz_cases/join_fuzz.rs).  Outputs are compared as sorted multisets of rows."""
import json
import os

import numpy as np
import pyarrow as pa
import pytest

import oracle
from gpu_helpers import gpu_ctx, random_keys
from util import assert_rows_equal, golden_filter, rows_of

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
JOINS = json.load(open(os.path.join(G, "hash_join.json")))
JT = {"Inner": "inner", "Left": "left", "Right": "right", "Full": "full", "LeftSemi": "left_semi",
      "RightSemi": "right_semi", "LeftAnti": "left_anti", "RightAnti": "right_anti"}
ALL_TYPES = list(oracle.JOIN_TYPES)


@pytest.fixture(scope="module")
def ctx():
    c = gpu_ctx()
    yield c
    c.close()


@pytest.fixture(autouse=True, params=["index", "general"])
def probe_path(request, monkeypatch):
    """Every case runs through both probe paths: the fused probe-side index (64-bit key images, join_fast.inc)
    and the general GroupTable path it falls back to for wide / string / dictionary keys."""
    if request.param == "general":
        monkeypatch.setenv("DFGPU_DISABLE_JOIN_INDEX", "1")
    else:
        monkeypatch.delenv("DFGPU_DISABLE_JOIN_INDEX", raising=False)
    return request.param


def gpu_join(ctx, jt, left, right, on, null_equals_null=False, left_schema=None, right_schema=None, filter=None):
    import datafusion_upstream_b200 as df
    ls = left_schema or left[0].schema
    rs = right_schema or right[0].schema
    j = df.GpuHashJoinExec(ctx, ls, rs, on, df.JoinType(oracle.JOIN_TYPES[jt]), df.PartitionMode.CollectLeft, null_equals_null)
    if filter is not None:
        j.set_filter(*filter)
    out = j.execute(left, right)
    names = j.schema.names
    j.close()
    return out, names


def batches(side):
    return [pa.record_batch([pa.array(c, pa.int32()) for c in b["columns"]], names=b["names"]) for b in side["batches"]]


@pytest.mark.parametrize("case", JOINS, ids=[c["name"] for c in JOINS])
def test_join_golden(ctx, case):
    left, right = batches(case["left"]), batches(case["right"])
    on = [(left[0].schema.names.index(l), right[0].schema.names.index(r)) for l, r in case["on"]]
    c = gpu_ctx(force_hash_collisions=True) if case.get("force_collisions") else ctx
    out, names = gpu_join(c, JT[case["join_type"]], left, right, on, filter=golden_filter(case))   # incl. the join_*_with_filter tables
    assert names == case["expected_header"]
    assert_rows_equal(rows_of(out), [tuple(r) for r in case["expected"]], what=case["name"])


def rand_side(rng, n, key_kinds, card, nulls, tag):
    cols = [random_keys(rng, n, k, card, nulls) for k in key_kinds]
    cols.append(pa.array(np.arange(n) + (1000000 if tag == "r" else 0), pa.int64()))
    cols.append(pa.array([f"{tag}{i % 13}" if i % 7 else None for i in range(n)], pa.string()))
    return pa.record_batch(cols, names=[f"{tag}k{i}" for i in range(len(key_kinds))] + [tag + "id", tag + "s"])


KEYS = [["int64"], ["int32", "int32"], ["utf8"], ["utf8", "int16"], ["date32"], ["decimal"], ["int64", "date32", "int32"], ["bool", "uint8"]]


@pytest.mark.parametrize("jt", ALL_TYPES)
@pytest.mark.parametrize("keys", KEYS, ids=lambda k: "+".join(k))
def test_join_fuzz_vs_oracle(ctx, jt, keys):
    rng = np.random.default_rng(sum(map(ord, jt + "".join(keys))))
    left = [rand_side(rng, n, keys, 60, 0.1, "l") for n in (150, 1, 90)]
    right = [rand_side(rng, n, keys, 75, 0.1, "r") for n in (400, 0, 333)]
    on = [(i, i) for i in range(len(keys))]
    out, _ = gpu_join(ctx, jt, left, right, on)
    want = oracle.hash_join(jt, left, right, on)
    assert_rows_equal(rows_of(out), rows_of(want), what=f"{jt} {keys}")


@pytest.mark.parametrize("jt", ALL_TYPES)
@pytest.mark.parametrize("keys", [["int64"], ["utf8", "int16"]], ids=lambda k: "+".join(k))
def test_join_filter_fuzz_vs_oracle(ctx, jt, keys):
    """Residual JoinFilter `l.id % 3 != r.id % 4 or r.s is null` (a predicate that is NULL-free but selective) through
    probe_candidates / probe_commit, all 8 join types, several build and probe batches incl. an empty one; a second predicate
    that evaluates to NULL for some pairs (NULL = not kept) and one that rejects everything."""
    import pyarrow.compute as pc
    rng = np.random.default_rng(sum(map(ord, jt + "".join(keys))) + 1)
    left = [rand_side(rng, n, keys, 25, 0.1, "l") for n in (150, 1, 90)]
    right = [rand_side(rng, n, keys, 30, 0.1, "r") for n in (400, 0, 333)]
    on = [(i, i) for i in range(len(keys))]
    nk = len(keys)
    cols = [("left", nk), ("right", nk), ("right", nk + 1)]          # l.id, r.id, r.s
    preds = [
        lambda b: pc.or_(pc.not_equal(pc.bit_wise_and(b.column(0), 3), pc.bit_wise_and(b.column(1), 3)), pc.is_null(b.column(2))),
        lambda b: pc.equal(b.column(2), pa.scalar("r3")),              # NULL where r.s is NULL
        lambda b: pa.array(np.zeros(b.num_rows, bool)),
    ]
    for i, pred in enumerate(preds):
        out, _ = gpu_join(ctx, jt, left, right, on, filter=(cols, pred))
        want = oracle.hash_join(jt, left, right, on, filter=(cols, pred))
        assert_rows_equal(rows_of(out), rows_of(want), what=f"{jt} {keys} predicate {i}")


def test_join_filter_protocol_errors(ctx):
    import datafusion_upstream_b200 as df
    l = pa.record_batch([pa.array([1, 2, 3], pa.int64())], names=["a"])
    r = pa.record_batch([pa.array([2, 3, 4], pa.int64())], names=["b"])
    j = df.GpuHashJoinExec(ctx, l.schema, r.schema, [("a", "b")], df.JoinType.Left)
    with pytest.raises(df.DfGpuError):
        j.set_filter([("left", 5)], lambda b: None)                    # column out of range: Plan error
    j.set_filter([("left", 0), ("right", 0)], None)
    j.build([l])
    with pytest.raises(df.DfGpuError) as e:
        j.probe(r)                                                     # dfgpu_join_probe refuses a filtered join
    assert "JoinFilter" in str(e.value)
    j.close()


@pytest.mark.parametrize("jt", ALL_TYPES)
def test_join_null_equals_null(ctx, jt):
    """INTERSECT/EXCEPT plans join with null_equals_null (hash_join.rs:1067-1076)."""
    rng = np.random.default_rng(4)
    left = [rand_side(rng, 120, ["int32", "utf8"], 12, 0.3, "l")]
    right = [rand_side(rng, 200, ["int32", "utf8"], 14, 0.3, "r")]
    on = [(0, 0), (1, 1)]
    out, _ = gpu_join(ctx, jt, left, right, on, null_equals_null=True)
    want = oracle.hash_join(jt, left, right, on, null_equals_null=True)
    assert_rows_equal(rows_of(out), rows_of(want), what=jt)


@pytest.mark.parametrize("jt", ALL_TYPES)
def test_join_empty_sides(ctx, jt):
    rng = np.random.default_rng(2)
    l, r = rand_side(rng, 50, ["int64"], 10, 0.0, "l"), rand_side(rng, 70, ["int64"], 10, 0.0, "r")
    for left, right in (([l.slice(0, 0)], [r]), ([l], [r.slice(0, 0)]), ([l.slice(0, 0)], [r.slice(0, 0)])):
        out, _ = gpu_join(ctx, jt, left, right, [(0, 0)])
        want = oracle.hash_join(jt, left, right, [(0, 0)])
        assert_rows_equal(rows_of(out), rows_of(want), what=jt)


def test_join_dictionary_keys(ctx):
    rng = np.random.default_rng(8)
    lk = random_keys(rng, 300, "dict_utf8", 40, 0.1)
    rk = random_keys(rng, 500, "utf8", 55, 0.1)
    left = [pa.record_batch([lk, pa.array(np.arange(300))], names=["k", "l"])]
    right = [pa.record_batch([rk.dictionary_encode(), pa.array(np.arange(500))], names=["k", "r"])]
    for jt in ("inner", "full", "left_anti"):
        out, _ = gpu_join(ctx, jt, left, right, [(0, 0)])
        want = oracle.hash_join(jt, left, right, [(0, 0)])
        assert_rows_equal(rows_of(out), rows_of(want), what=jt)


def test_join_q3_shape_large(ctx):
    """Q3 shape (config C3) at a size the oracle finishes in seconds: PK-FK joins, duplicates on the probe
    side, multi-chunk probe; plus a size-independent property (every probe row with an existing key appears)."""
    rng = np.random.default_rng(6)
    nb, np_ = 200_000, 1_500_000
    okey = rng.permutation(nb * 4)[:nb].astype(np.int64)
    orders = pa.record_batch([pa.array(okey), pa.array(rng.integers(8000, 10000, nb).astype(np.int32), pa.int32()).cast(pa.date32()),
                              pa.array(np.zeros(nb, np.int32))], names=["o_orderkey", "o_orderdate", "o_shippriority"])
    lkey = rng.integers(0, nb * 4, np_).astype(np.int64)
    line = pa.record_batch([pa.array(lkey), pa.array(rng.integers(0, 10**7, np_), pa.int64())], names=["l_orderkey", "l_price"])
    out, _ = gpu_join(ctx, "inner", [orders], [line.slice(0, 700_000), line.slice(700_000)], [(0, 0)])
    got = pa.Table.from_batches(out)
    want = oracle.hash_join("inner", [orders], [line], [(0, 0)])
    assert got.num_rows == want.num_rows == int(np.isin(lkey, okey).sum())
    assert_rows_equal(rows_of(got), rows_of(want))


@pytest.mark.parametrize("jt", ["inner", "left", "full"])
def test_join_large_build_packed_rows_and_block_build(ctx, jt):
    """A build side of > 2^20 rows: the index is built block by block from partitioned rows (more blocks than one CTA histogram
    holds in the Partitioned variant below is covered by bench sizes), keys that repeat get their runs in rows[], and the build
    columns that reach the output are gathered from row-wise packed rows (Inner / Left; Full keeps the column-wise gather because
    unmatched probe rows carry NULL build columns) - mixed widths, one nullable and one string column that stay column-wise."""
    rng = np.random.default_rng(11)
    nb, np_ = 1_200_000, 120_000
    bk = rng.integers(0, 900_000, nb).astype(np.int64)          # ~ 25 % of the keys repeat
    bk[:5] = [-1, -1, 2**63 - 1, -2**63, 7]
    build = pa.record_batch([
        pa.array(bk),
        pa.array(rng.integers(-2**31, 2**31, nb), pa.int32()),
        pa.array(rng.integers(0, 2**15, nb), pa.int16()),
        pa.array(rng.random(nb), pa.float64()),
        pa.array(rng.integers(-10**9, 10**9, nb), pa.int64()).cast(pa.decimal128(25, 2)),
        pa.array(rng.integers(0, 100, nb), pa.int64(), mask=rng.random(nb) < 0.05),
        pa.array([f"s{v}" for v in rng.integers(0, 1000, nb)]),
        pa.array(rng.integers(0, 255, nb), pa.uint8()),
    ], names=["k", "i32", "i16", "f64", "dec", "nullable", "str", "u8"])
    pk = rng.integers(-100_000, 1_000_000, np_).astype(np.int64)
    pk[:3] = [-1, 2**63 - 1, 12345678901]
    probe = pa.record_batch([pa.array(pk), pa.array(np.arange(np_), pa.int64())], names=["pk", "y"])
    out, _ = gpu_join(ctx, jt, [build.slice(0, 500_000), build.slice(500_000)], [probe.slice(0, 150_000), probe.slice(150_000)], [(0, 0)])
    want = oracle.hash_join(jt, [build], [probe], [(0, 0)])
    assert_rows_equal(rows_of(out), rows_of(want), what=jt)


@pytest.mark.parametrize("jt", ALL_TYPES)
@pytest.mark.parametrize("n_probe", [1, 1023, 1024, 1025, 2047, 2048, 2049, 3 * 2048 + 5])
def test_join_index_edges(ctx, jt, n_probe):
    """Fused-probe edge cases: tile boundaries (1024 rows per CTA), build keys with 1 / many rows, the key whose
    image is all ones (-1) and, under null_equals_null, the NULL key of a single 8-byte column (both live in
    side entries of the index)."""
    rng = np.random.default_rng(n_probe)
    bk = np.concatenate([np.arange(40), np.repeat([3, 7, -1], [5, 300, 2]), [2**63 - 1, -2**63]]).astype(np.int64)
    bmask = np.zeros(len(bk), bool)
    bmask[[1, 50]] = True
    left = [pa.record_batch([pa.array(bk, pa.int64(), mask=bmask), pa.array(np.arange(len(bk)), pa.int32())], names=["a", "x"])]
    pk = rng.choice(np.array([-1, 3, 7, 11, 39, 40, 999, 2**63 - 1, -2**63], dtype=np.int64), n_probe)
    pmask = rng.random(n_probe) < 0.1
    right = [pa.record_batch([pa.array(pk, pa.int64(), mask=pmask), pa.array(np.arange(n_probe), pa.int64())], names=["b", "y"])]
    for nen in (False, True):
        out, _ = gpu_join(ctx, jt, left, right, [(0, 0)], null_equals_null=nen)
        want = oracle.hash_join(jt, left, right, [(0, 0)], null_equals_null=nen)
        assert_rows_equal(rows_of(out), rows_of(want), what=f"{jt} n={n_probe} null_equals_null={nen}")


def test_join_errors(ctx):
    import datafusion_upstream_b200 as df
    l = pa.record_batch([pa.array([1], pa.int64())], names=["a"])
    r = pa.record_batch([pa.array(["x"])], names=["a"])
    with pytest.raises(df.DfGpuError) as e:
        gpu_join(ctx, "inner", [l], [r], [(0, 0)])
    assert e.value.kind == "Plan"
    with pytest.raises(df.DfGpuError) as e:
        gpu_join(ctx, "inner", [l], [l], [])
    assert e.value.kind == "Plan" and "non-empty" in str(e.value)

"""Parity of the partition path of the two-level aggregation (fastagg_part.inc: sample -> hot keys, two-pass radix
partition of the bare key images, per-partition count) with the oracle.  The library takes this path by itself only for
COUNT(*)-shaped operators over >= 4 Mi rows whose sample shows more distinct keys than a level-1 table holds; the
DFGPU_FA_PART hook forces it on small inputs so every branch (hot keys, side groups, overflow rounds, several pushes,
compaction after 8 pieces, pushes that fall back to the cache path) runs under pytest."""
import numpy as np
import pyarrow as pa
import pytest

from gpu_helpers import gpu_agg, gpu_ctx, oracle_agg
from util import assert_tables_equal

pytestmark = pytest.mark.gpu

COUNT = [("count", None, "COUNT(*)", None)]


def kernels_of(c, fn):
    c.set_profiling(True)
    try:
        out = fn()
        names = {r["name"] for r in c.profile()}
    finally:
        c.set_profiling(False)
    return out, names


def zipf_keys(rng, n, universe, kind="int64"):
    r = np.exp(rng.random(n) * np.log(universe)).astype(np.int64)
    with np.errstate(over="ignore"):
        k = (r.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)).view(np.int64)
    return k if kind == "int64" else (r % 2_000_000).astype(np.int32)


@pytest.mark.parametrize("s2max", [None, "64"], ids=["default", "tiny_tables"])
@pytest.mark.parametrize("shape", ["zipf", "uniform", "few", "nulls_and_minus_one"])
def test_forced_partition_path_vs_oracle(monkeypatch, shape, s2max):
    monkeypatch.setenv("DFGPU_FA_PART", "1")
    if s2max:
        monkeypatch.setenv("DFGPU_FA_S2MAX", s2max)     # partitions overflow their tables: record rounds on the next hash bits
    c = gpu_ctx()
    try:
        rng = np.random.default_rng(len(shape))
        n = 300_007
        if shape == "zipf":
            keys = pa.array(zipf_keys(rng, n, 400_000), pa.int64())
        elif shape == "uniform":
            keys = pa.array(rng.integers(-2**62, 2**62, n), pa.int64())
        elif shape == "few":
            keys = pa.array(rng.integers(0, 7, n), pa.int64())               # everything is a hot key
        else:
            raw = zipf_keys(rng, n, 50_000)
            raw[rng.random(n) < 0.05] = -1                                    # the all-ones image: side group 1
            keys = pa.array(raw, pa.int64(), mask=rng.random(n) < 0.07)       # NULL key: side group 0
        b = pa.record_batch([keys], names=["k"])
        got, names = kernels_of(c, lambda: gpu_agg(c, "single", [b], ["k"], COUNT))
        assert {"k_hp_hist", "k_hp_part1", "k_hp_part2", "k_hp_merge"} <= names, names
        want = oracle_agg("single", [b], ["k"], COUNT)
        assert_tables_equal(got, want, key_cols=[0], what=f"{shape} s2max={s2max}")
    finally:
        c.close()


def test_packed_keys_and_count_of_a_column_without_nulls(monkeypatch):
    """A multi-column key image (Int32, Int16 NOT NULL + a nullable Bool) and COUNT(x) over a column without a validity bitmap
    (= COUNT(*)) take the path too; COUNT(x) over a column WITH nulls does not (its rows are not all counted)."""
    monkeypatch.setenv("DFGPU_FA_PART", "1")
    c = gpu_ctx()
    try:
        rng = np.random.default_rng(5)
        n = 200_000
        b = pa.record_batch([pa.array(zipf_keys(rng, n, 300_000, "int32"), pa.int32()), pa.array(rng.integers(0, 50, n).astype(np.int16), pa.int16()),
                             pa.array(rng.random(n) < 0.5, pa.bool_(), mask=rng.random(n) < 0.1), pa.array(rng.integers(0, 9, n), pa.int64()),
                             pa.array(rng.integers(0, 9, n), pa.int64(), mask=rng.random(n) < 0.3)], names=["a", "b", "f", "x", "xn"])
        spec = [("count", "x", "COUNT(x)", None)]
        got, names = kernels_of(c, lambda: gpu_agg(c, "single", [b], ["a", "b", "f"], spec))
        assert "k_hp_merge" in names, names
        assert_tables_equal(got, oracle_agg("single", [b], ["a", "b", "f"], spec), key_cols=[0, 1, 2])
        spec = [("count", "xn", "COUNT(xn)", None)]
        got, names = kernels_of(c, lambda: gpu_agg(c, "single", [b], ["a", "b", "f"], spec))
        assert "k_hp_merge" not in names and "k_fa_l1" in names, names
        assert_tables_equal(got, oracle_agg("single", [b], ["a", "b", "f"], spec), key_cols=[0, 1, 2])
    finally:
        c.close()


def test_many_pushes_compaction_and_mixed_paths(monkeypatch):
    """11 pushes: after 8 pieces the operator folds them into partial-state records and goes on; one push carries a FILTER-less
    COUNT(*) over a batch whose key column has a validity bitmap (still eligible) - and num_groups() mid-stream compacts too."""
    monkeypatch.setenv("DFGPU_FA_PART", "1")
    import datafusion_upstream_b200 as df
    c = gpu_ctx(coalesce_rows=-1)
    try:
        rng = np.random.default_rng(11)
        batches = []
        for i in range(11):
            n = 40_000 + 1000 * i
            k = zipf_keys(rng, n, 100_000)
            batches.append(pa.record_batch([pa.array(k, pa.int64(), mask=(rng.random(n) < 0.02) if i % 3 == 0 else None)], names=["k"]))
        op = df.GpuAggregateExec(c, df.AggregateMode.Single, batches[0].schema, ["k"], [df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")])
        for i, b in enumerate(batches):
            op.push(b)
            if i == 4:
                want_so_far = oracle_agg("single", batches[:5], ["k"], COUNT).num_rows
                assert op.num_groups() == want_so_far
        got = op.emit()
        op.close()
        assert_tables_equal(got, oracle_agg("single", batches, ["k"], COUNT), key_cols=[0])
    finally:
        c.close()


def test_partial_mode_state_and_final_merge(monkeypatch):
    """Partial mode through the partition path emits COUNT(*)[count] states the reference's Final merges (count.rs:135-160)."""
    monkeypatch.setenv("DFGPU_FA_PART", "1")
    c = gpu_ctx()
    try:
        rng = np.random.default_rng(2)
        parts = [[pa.record_batch([pa.array(zipf_keys(rng, 150_000, 60_000), pa.int64())], names=["k"])] for _ in range(3)]
        states = [gpu_agg(c, "partial", p, ["k"], COUNT) for p in parts]
        for s, p in zip(states, parts):
            assert_tables_equal(s, oracle_agg("partial", p, ["k"], COUNT), key_cols=[0], what="partial state")
        monkeypatch.delenv("DFGPU_FA_PART")
        fin = [("count", None, "COUNT(*)", None)]
        got = gpu_agg(c, "final_partitioned", states, [0], fin)
        assert_tables_equal(got, oracle_agg("single", [b for p in parts for b in p], ["k"], COUNT), key_cols=[0])
    finally:
        c.close()


def test_library_picks_the_path_by_sample():
    """No hook: 6 M rows over ~1.5 M distinct keys -> partition path; the same rows over 1000 keys -> cache path; both exact."""
    c = gpu_ctx()
    try:
        rng = np.random.default_rng(8)
        n = 6_000_000
        for universe, expect in ((3_000_000, True), (1000, False)):
            k = zipf_keys(rng, n, universe)
            b = pa.record_batch([pa.array(k, pa.int64())], names=["k"])
            got, names = kernels_of(c, lambda: gpu_agg(c, "single", [b], ["k"], COUNT))
            assert ("k_hp_merge" in names) == expect, (universe, names)
            uk, uc = np.unique(k, return_counts=True)
            gk, gc = got.column(0).to_numpy(), got.column(1).to_numpy()
            o = np.argsort(gk)
            assert (gk[o] == uk).all() and (gc[o] == uc).all()
    finally:
        c.close()

"""Shared helpers for the parity tests: canonical sorted-multiset comparison."""
import decimal
import math

import pyarrow as pa


def rows_of(tbl):
    """RecordBatch/Table/list of batches -> list of row tuples (dictionaries decoded)."""
    if isinstance(tbl, (list, tuple)):
        out = []
        for b in tbl:
            if b is not None:
                out += rows_of(b)
        return out
    cols = []
    for c in tbl.columns:
        if isinstance(c, pa.ChunkedArray):
            c = c.combine_chunks()
        if pa.types.is_dictionary(c.type):
            c = c.cast(c.type.value_type)
        cols.append(c.to_pylist())
    return list(zip(*cols)) if cols else []


def _key(v):
    if v is None:
        return (0, 0)
    if isinstance(v, float):
        if math.isnan(v):
            return (2, 0)
        return (1, v)
    if isinstance(v, (bytes, str)):
        return (1, v)
    if isinstance(v, decimal.Decimal):
        return (1, v)
    if isinstance(v, bool):
        return (1, int(v))
    try:
        return (1, v)
    except Exception:
        return (1, str(v))


def sort_rows(rows, key_cols=None):
    n = len(rows[0]) if rows else 0
    idx = list(range(n)) if key_cols is None else list(key_cols)
    return sorted(rows, key=lambda r: tuple(_key(r[i]) for i in idx))


def values_equal(a, b, rel=1e-9):
    if a is None or b is None:
        return a is None and b is None
    if isinstance(a, float) or isinstance(b, float):
        a, b = float(a), float(b)
        if math.isnan(a) or math.isnan(b):
            return math.isnan(a) and math.isnan(b)
        if a == b:
            return True
        return abs(a - b) <= rel * max(abs(a), abs(b))   # north_star: Float64 within 1e-9 relative
    return a == b                                          # bit-exact for everything else


def assert_rows_equal(got, want, key_cols=None, rel=1e-9, ordered=False, what=""):
    if not ordered:
        got, want = sort_rows(got, key_cols), sort_rows(want, key_cols)
    assert len(got) == len(want), f"{what}: row count {len(got)} != {len(want)}"
    for i, (g, w) in enumerate(zip(got, want)):
        assert len(g) == len(w), f"{what}: width {len(g)} != {len(w)}"
        for j, (x, y) in enumerate(zip(g, w)):
            assert values_equal(x, y, rel), f"{what}: row {i} col {j}: {x!r} != {y!r}\n got={g}\nwant={w}"


def assert_tables_equal(got, want, key_cols=None, rel=1e-9, ordered=False, check_types=True, what=""):
    if check_types and not isinstance(got, (list, tuple)) and not isinstance(want, (list, tuple)):
        gt = [f.type for f in got.schema]
        wt = [f.type for f in want.schema]
        assert gt == wt, f"{what}: schema types {gt} != {wt}"
        assert got.schema.names == want.schema.names, f"{what}: names {got.schema.names} != {want.schema.names}"
    assert_rows_equal(rows_of(got), rows_of(want), key_cols, rel, ordered, what)


def golden_filter(case):
    """The JoinFilter of a golden join case as (column_indices, predicate over the intermediate batch), or None:
    `column <op> literal | column` over Int32 (hash_join.rs `prepare_join_filter` and the *_with_filter tests)."""
    import pyarrow as pa
    import pyarrow.compute as pc
    f = case.get("filter")
    if not f:
        return None
    ops = {"Gt": pc.greater, "NotEq": pc.not_equal, "Lt": pc.less, "Eq": pc.equal, "GtEq": pc.greater_equal, "LtEq": pc.less_equal}
    e = f["expr"]

    def predicate(batch):
        rhs = pa.scalar(e["rhs"]["lit"], pa.int32()) if "lit" in e["rhs"] else batch.column(e["rhs"]["col"])
        return ops[e["op"]](batch.column(e["lhs"]), rhs)
    return [(s, i) for s, i in f["column_indices"]], predicate

"""CPU-side checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and exports
every symbol include/dfgpu.h declares; without a CUDA device it fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, "datafusion_upstream_b200", "csrc")], check=True)
    from datafusion_upstream_b200 import _native
    return _native.lib()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "dfgpu.h")).read()
    declared = set(re.findall(r"\b(dfgpu_[a-z0-9_]+)\s*\(", hdr))
    from datafusion_upstream_b200 import _native
    assert declared == set(_native.SYMBOLS), declared ^ set(_native.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s


def test_sm100a_only():
    so = os.path.join(ROOT, "datafusion_upstream_b200", "libdfgpu.so")
    out = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from datafusion_upstream_b200 import _native
    h = C.c_void_p()
    st = lib.dfgpu_ctx_create(None, C.byref(h))
    assert st == 6 and b"no CPU fallback" in lib.dfgpu_last_error()
    import datafusion_upstream_b200 as df
    with pytest.raises(df.DfGpuError):
        df.GpuContext()


def test_product_does_not_touch_oracle():
    pkg = os.path.join(ROOT, "datafusion_upstream_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".inc", ".cpp")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "liboracle" not in src and "df_oracle" not in src, f


def test_schema_roundtrip_without_gpu(lib):
    """Arrow C schema export of the operator output schema is pure host code."""
    import pyarrow as pa
    from datafusion_upstream_b200 import _native as N
    s = N.export_schema(pa.schema([("a", pa.int64()), ("d", pa.decimal128(15, 2)), ("s", pa.dictionary(pa.int32(), pa.string()))]))
    assert s.format == b"+s" and s.n_children == 3
    N.release_schema(s)


def test_struct_layouts_match_the_header(tmp_path):
    """include/dfgpu.h compiles as plain C (what cgo / bindgen / cffi would consume) and the ctypes mirrors of its structs
    have the same size and field offsets - an appended dfgpu_options field that the binding misses would otherwise pass
    silently (the library reads past the caller's struct)."""
    from datafusion_upstream_b200 import _native as N
    structs = {"dfgpu_options": N.Options, "dfgpu_agg_desc": N.AggDesc, "dfgpu_metrics": N.Metrics, "dfgpu_expr_node": N.ExprNode,
               "struct ArrowArray": N.ArrowArray, "struct ArrowSchema": N.ArrowSchema}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "dfgpu.h"', 'int main(void) {']
    for cname, ct in structs.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in ct._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = str(tmp_path / "layout")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", exe], check=True)
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout
    got = {}
    for ln in out.strip().splitlines():
        name, field, v = ln.rsplit(" ", 2)
        got[(name, field)] = int(v)
    for cname, ct in structs.items():
        assert got[(cname, "size")] == C.sizeof(ct), cname
        for fname, _ in ct._fields_:
            assert got[(cname, fname)] == getattr(ct, fname).offset, (cname, fname)

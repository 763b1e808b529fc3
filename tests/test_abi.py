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

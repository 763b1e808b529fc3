"""ctypes binding of ``libdfgpu.so`` (the C ABI in ``include/dfgpu.h``).

This is the same binding a Rust maintainer would write with ``extern "C"`` + ``arrow::ffi``
(see INTEGRATION.md); Python is only the test / bench driver.  There is no CPU fallback: importing
this module without the built library, or creating a context without a CUDA device, fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import pyarrow as pa

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdfgpu.so")


class DfGpuError(RuntimeError):
    """Mirrors DataFusionError; ``code`` is the dfgpu_status."""

    NAMES = {1: "Internal", 2: "Plan", 3: "Execution", 4: "ResourcesExhausted", 5: "NotImplemented", 6: "Cuda", 7: "Nccl"}

    def __init__(self, code: int, msg: str):
        super().__init__(f"{self.NAMES.get(code, code)}: {msg}")
        self.code = code
        self.kind = self.NAMES.get(code, str(code))


class ArrowSchema(C.Structure):
    pass


class ArrowArray(C.Structure):
    pass


ArrowSchema._fields_ = [
    ("format", C.c_char_p), ("name", C.c_char_p), ("metadata", C.c_char_p), ("flags", C.c_int64),
    ("n_children", C.c_int64), ("children", C.POINTER(C.POINTER(ArrowSchema))),
    ("dictionary", C.POINTER(ArrowSchema)), ("release", C.c_void_p), ("private_data", C.c_void_p),
]
ArrowArray._fields_ = [
    ("length", C.c_int64), ("null_count", C.c_int64), ("offset", C.c_int64), ("n_buffers", C.c_int64),
    ("n_children", C.c_int64), ("buffers", C.POINTER(C.c_void_p)), ("children", C.POINTER(C.POINTER(ArrowArray))),
    ("dictionary", C.POINTER(ArrowArray)), ("release", C.c_void_p), ("private_data", C.c_void_p),
]


class Options(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("batch_size", C.c_int32), ("memory_limit", C.c_int64), ("load_factor", C.c_double),
        ("force_hash_collisions", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
        ("nccl_unique_id", C.c_void_p), ("chunk_rows", C.c_int64), ("coalesce_rows", C.c_int64),
    ]


class ExprNode(C.Structure):
    _fields_ = [("kind", C.c_int32), ("column", C.c_int32), ("left", C.c_int32), ("right", C.c_int32),
                ("lit_lo", C.c_uint64), ("lit_hi", C.c_uint64), ("lit_precision", C.c_int32), ("lit_scale", C.c_int32)]


class AggDesc(C.Structure):
    _fields_ = [("func", C.c_int32), ("name", C.c_char_p), ("arg_col", C.c_int32), ("filter_col", C.c_int32)]


class Metrics(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "elapsed_compute_ns", "output_rows", "output_batches", "input_rows", "input_batches", "build_input_rows",
        "build_input_batches", "build_mem_used", "build_time_ns", "join_time_ns", "h2d_bytes", "d2h_bytes",
        "kernel_launches", "nvlink_bytes", "device_mem_used")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# every symbol include/dfgpu.h declares (tests check the library exports all of them)
SYMBOLS = [
    "dfgpu_last_error", "dfgpu_version", "dfgpu_options_default", "dfgpu_ctx_create", "dfgpu_ctx_destroy",
    "dfgpu_ctx_synchronize", "dfgpu_ctx_set_profiling", "dfgpu_ctx_profile_json", "dfgpu_ctx_stream", "dfgpu_ctx_kernel_launches", "dfgpu_host_alloc", "dfgpu_host_free",
    "dfgpu_batch_import", "dfgpu_batch_export", "dfgpu_batch_num_rows", "dfgpu_batch_num_columns", "dfgpu_batch_schema",
    "dfgpu_batch_concat", "dfgpu_batch_slice", "dfgpu_batch_free", "dfgpu_create_hashes", "dfgpu_agg_create",
    "dfgpu_agg_output_schema", "dfgpu_agg_push", "dfgpu_agg_push_arrow", "dfgpu_agg_push_arrow_batches", "dfgpu_agg_num_groups", "dfgpu_agg_emit",
    "dfgpu_agg_emit_arrow", "dfgpu_agg_metrics", "dfgpu_agg_destroy", "dfgpu_join_create", "dfgpu_join_output_schema",
    "dfgpu_join_build_push", "dfgpu_join_build_finish", "dfgpu_join_probe", "dfgpu_join_emit_unmatched",
    "dfgpu_join_metrics", "dfgpu_join_destroy", "dfgpu_partition_hash", "dfgpu_exchange_all_to_all",
    "dfgpu_nccl_get_unique_id", "dfgpu_exchange_all_gather", "dfgpu_batch_project", "dfgpu_agg_exchange", "dfgpu_batch_filter", "dfgpu_batch_eval_decimal",
    "dfgpu_coalescer_create", "dfgpu_coalescer_push", "dfgpu_coalescer_finish", "dfgpu_coalescer_destroy",
    "dfgpu_partition_i64_noalloc", "dfgpu_agg_exchange_virtual",
    "dfgpu_agg_emit_first", "dfgpu_hash_contract", "dfgpu_join_set_filter", "dfgpu_join_probe_candidates", "dfgpu_join_probe_commit",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  datafusion_upstream_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    P = C.c_void_p
    L.dfgpu_last_error.restype = C.c_char_p
    L.dfgpu_version.restype = C.c_char_p
    L.dfgpu_hash_contract.restype = C.c_int32
    L.dfgpu_options_default.argtypes = [C.POINTER(Options)]
    L.dfgpu_ctx_create.argtypes = [C.POINTER(Options), C.POINTER(P)]
    L.dfgpu_ctx_destroy.argtypes = [P]
    L.dfgpu_ctx_synchronize.argtypes = [P]
    L.dfgpu_ctx_set_profiling.argtypes = [P, C.c_int32]
    L.dfgpu_ctx_profile_json.argtypes = [P, C.c_char_p, C.c_int64]
    L.dfgpu_ctx_stream.restype = P
    L.dfgpu_ctx_stream.argtypes = [P]
    L.dfgpu_ctx_kernel_launches.restype = C.c_int64
    L.dfgpu_ctx_kernel_launches.argtypes = [P]
    L.dfgpu_host_alloc.argtypes = [P, C.c_int64, C.POINTER(P)]
    L.dfgpu_host_free.argtypes = [P, P]
    L.dfgpu_batch_import.argtypes = [P, P, P, C.c_int32, C.POINTER(P)]
    L.dfgpu_batch_export.argtypes = [P, P, P, P]
    L.dfgpu_batch_num_rows.restype = C.c_int64
    L.dfgpu_batch_num_rows.argtypes = [P]
    L.dfgpu_batch_num_columns.argtypes = [P]
    L.dfgpu_batch_schema.argtypes = [P, P]
    L.dfgpu_batch_concat.argtypes = [P, C.POINTER(P), C.c_int32, C.POINTER(P)]
    L.dfgpu_batch_slice.argtypes = [P, P, C.c_int64, C.c_int64, C.POINTER(P)]
    L.dfgpu_agg_exchange.argtypes = [P]
    L.dfgpu_agg_exchange_virtual.argtypes = [C.POINTER(P), C.c_int32]
    L.dfgpu_batch_project.argtypes = [P, P, C.POINTER(C.c_int32), C.c_int32, C.POINTER(P)]
    L.dfgpu_batch_filter.argtypes = [P, P, C.c_int32, C.POINTER(P)]
    L.dfgpu_batch_eval_decimal.argtypes = [P, P, C.POINTER(ExprNode), C.c_int32, C.c_char_p, C.POINTER(P)]
    L.dfgpu_coalescer_create.argtypes = [P, P, C.c_int64, C.POINTER(P)]
    L.dfgpu_coalescer_push.argtypes = [P, P, C.POINTER(P)]
    L.dfgpu_coalescer_finish.argtypes = [P, C.POINTER(P)]
    L.dfgpu_coalescer_destroy.argtypes = [P]
    L.dfgpu_coalescer_destroy.restype = None
    L.dfgpu_batch_free.argtypes = [P]
    L.dfgpu_create_hashes.argtypes = [P, P, C.POINTER(C.c_int32), C.c_int32, C.c_uint64, P]
    L.dfgpu_agg_create.argtypes = [P, C.c_int32, P, C.POINTER(C.c_int32), C.c_int32, C.POINTER(AggDesc), C.c_int32, C.POINTER(P)]
    L.dfgpu_agg_output_schema.argtypes = [P, P]
    L.dfgpu_agg_push.argtypes = [P, P]
    L.dfgpu_agg_push_arrow.argtypes = [P, P]
    L.dfgpu_agg_push_arrow_batches.argtypes = [P, C.POINTER(P), C.c_int32]
    L.dfgpu_agg_num_groups.argtypes = [P, C.POINTER(C.c_int64)]
    L.dfgpu_agg_emit.argtypes = [P, C.POINTER(P)]
    L.dfgpu_agg_emit_arrow.argtypes = [P, P, P]
    L.dfgpu_agg_emit_first.argtypes = [P, C.c_int64, C.POINTER(P)]
    L.dfgpu_agg_metrics.argtypes = [P, C.POINTER(Metrics)]
    L.dfgpu_agg_destroy.argtypes = [P]
    L.dfgpu_join_create.argtypes = [P, C.c_int32, C.c_int32, C.c_int32, P, P, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                    C.c_int32, C.POINTER(P)]
    L.dfgpu_join_output_schema.argtypes = [P, P]
    L.dfgpu_join_build_push.argtypes = [P, P]
    L.dfgpu_join_build_finish.argtypes = [P]
    L.dfgpu_join_probe.argtypes = [P, P, C.POINTER(P)]
    L.dfgpu_join_emit_unmatched.argtypes = [P, C.POINTER(P)]
    L.dfgpu_join_set_filter.argtypes = [P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int32]
    L.dfgpu_join_probe_candidates.argtypes = [P, P, C.POINTER(P)]
    L.dfgpu_join_probe_commit.argtypes = [P, P, C.c_int32, C.POINTER(P)]
    L.dfgpu_join_metrics.argtypes = [P, C.POINTER(Metrics)]
    L.dfgpu_join_destroy.argtypes = [P]
    L.dfgpu_partition_hash.argtypes = [P, P, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.POINTER(P)]
    L.dfgpu_exchange_all_to_all.argtypes = [P, C.POINTER(P), P, C.POINTER(P)]
    L.dfgpu_exchange_all_gather.argtypes = [P, P, P, C.POINTER(P)]
    L.dfgpu_nccl_get_unique_id.argtypes = [P]
    L.dfgpu_partition_i64_noalloc.argtypes = [P, P, P, C.c_int64, C.c_int32, P, P]
    _lib = L
    return L


def check(status: int):
    if status != 0:
        raise DfGpuError(status, (lib().dfgpu_last_error() or b"").decode())


def export_schema(schema: pa.Schema):
    """pa.Schema -> (ArrowSchema struct, keepalive)."""
    s = ArrowSchema()
    schema._export_to_c(C.addressof(s))
    return s


def release_schema(s: ArrowSchema):
    if s.release:
        C.CFUNCTYPE(None, C.POINTER(ArrowSchema))(s.release)(C.byref(s))


def import_schema(s: ArrowSchema) -> pa.Schema:
    return pa.Schema._import_from_c(C.addressof(s))

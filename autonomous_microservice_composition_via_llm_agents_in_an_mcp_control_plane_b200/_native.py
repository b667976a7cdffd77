"""ctypes binding of ``libmcp_b200.so`` (include/mcp_b200.h).

The library is the product; there is no Python or CPU stand-in for it.  If it is missing it is
built once with nvcc; if that fails, importing callers get a loud ``RuntimeError``.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

from . import build as _build

LIB_PATH: Path = _build.OUT
_lib = None

METRIC_COSINE, METRIC_L2 = 0, 1
PATH_AUTO, PATH_STREAM_FP32, PATH_GEMM_TF32, PATH_EXACT_FP64, PATH_GEMM_F16 = 0, 1, 2, 3, 4
MAX_K = 256
MAX_DAG_NODES = 64
MAX_WIDE_DAG_NODES = 255
NOALLOC_MAX_ROWS, NOALLOC_MAX_QUERIES = 1 << 20, 64


class McpStats(C.Structure):
    _fields_ = [
        ("searches", C.c_int64), ("queries", C.c_int64), ("rows_scanned", C.c_int64),
        ("bytes_scanned", C.c_int64), ("candidates_rescored", C.c_int64),
        ("certificate_failures", C.c_int64), ("exact_scans", C.c_int64),
        ("kernel_launches", C.c_int64), ("coarse_timed", C.c_int64), ("coarse_ms", C.c_double),
        ("last_path", C.c_int32), ("last_kprime", C.c_int32),
    ]


class McpError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"mcp_b200 error {code}: {msg}")
        self.code = code


_vp, _i32, _i64 = C.c_void_p, C.c_int, C.c_int64
# name -> (restype, argtypes); every symbol include/mcp_b200.h declares
SIGNATURES = {
    "mcp_index_create": (_i32, [C.POINTER(_vp), _i32, _i32, _i32]),
    "mcp_index_create_multi": (_i32, [C.POINTER(_vp), C.POINTER(C.c_int), _i32, _i32, _i32]),
    "mcp_index_destroy": (_i32, [_vp]),
    "mcp_index_load": (_i32, [_vp, _vp, _i64, _i64]),
    "mcp_index_load_file": (_i32, [_vp, C.c_char_p, _i64, _i64]),
    "mcp_index_adopt_device": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp]),
    "mcp_index_update_rows": (_i32, [_vp, _i64, _vp, _i64]),
    "mcp_index_upsert_rows": (_i32, [_vp, _vp, _vp, _i64]),
    "mcp_index_set_penalty": (_i32, [_vp, _vp]),
    "mcp_index_set_penalty_device": (_i32, [_vp, _vp, _vp]),
    "mcp_index_search": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "mcp_index_search_device": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "mcp_index_repair_device": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "mcp_index_coarse_scores_device": (_i32, [_vp, _i32, _vp, _i32, _vp, _vp]),
    "mcp_index_search_exact_noalloc_device": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "mcp_index_sync_status": (_i32, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "mcp_index_set_option": (_i32, [_vp, C.c_char_p, _i64]),
    "mcp_index_get_stats": (_i32, [_vp, C.POINTER(McpStats)]),
    "mcp_index_rows": (_i64, [_vp]),
    "mcp_merge_topk_device": (_i32, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "mcp_merge_packed_block_bytes": (_i64, [_i32, _i32]),
    "mcp_merge_topk_packed_device": (_i32, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "mcp_dag_levels": (_i32, [_i32, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "mcp_dag_levels_device": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "mcp_dag_levels_wide": (_i32, [_i32, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "mcp_dag_levels_wide_device": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "mcp_dag_launch_count": (_i64, [_i32]),
    "mcp_last_error": (C.c_char_p, []),
    "mcp_device_count": (_i32, []),
    "mcp_build_info": (C.c_char_p, []),
}


def lib() -> C.CDLL:
    """Load (building first if needed) the CUDA library.  Never falls back to anything else."""
    global _lib
    if _lib is None:
        try:
            _build.build()
        except Exception as e:  # noqa: BLE001
            if not LIB_PATH.exists():
                raise RuntimeError(
                    f"mcp_b200: {LIB_PATH.name} is missing and could not be built ({e}). "
                    "This package has no CPU fallback.") from e
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so is stale: fail loudly
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise McpError(rc, (lib().mcp_last_error() or b"").decode("utf-8", "replace"))


def device_count() -> int:
    return int(lib().mcp_device_count())


def build_info() -> str:
    return lib().mcp_build_info().decode()

"""``ServiceIndex`` -- the HBM-resident image of ``service_schemas.input_schema_vector``.

Host face of ``mcp_index_*`` (include/mcp_b200.h).  It replaces what
``GraphPlanner.__init__`` / ``_fetch_embeddings_metadata`` do with a pgvector connection
(control_plane.py:46-55): rows are loaded once, then ``search`` answers
``ORDER BY input_schema_vector <op> :q (+ penalty) LIMIT k`` exactly.

Two faces of the same calls:
  * numpy in / numpy out (``load``, ``set_penalty``, ``search``): host buffers, copies inside;
  * torch CUDA tensors (``adopt``, ``search_device``): device pointers on the current stream,
    nothing copied -- torch is only the allocator / stream provider here.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _native as nat

_METRICS = {"cosine": nat.METRIC_COSINE, "l2": nat.METRIC_L2, 0: 0, 1: 1}
_PATHS = {"auto": nat.PATH_AUTO, "stream_fp32": nat.PATH_STREAM_FP32, "gemm_tf32": nat.PATH_GEMM_TF32,
          "exact_fp64": nat.PATH_EXACT_FP64, "gemm_f16": nat.PATH_GEMM_F16}


def _np_ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


def _stream_ptr() -> C.c_void_p:
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class ServiceIndex:
    def __init__(self, dim: int, metric="cosine", device=0):
        """``device``: one GPU ordinal, or a list of them -- then the table is row-sharded over those GPUs behind this one
        handle (``mcp_index_create_multi``): ``load`` / ``set_penalty`` / ``update_rows`` / ``search`` work as usual, the
        device-pointer calls (``adopt``, ``search_device`` ...) do not."""
        self._lib = nat.lib()
        self._h = C.c_void_p()
        self.dim = int(dim)
        self.metric = _METRICS[metric]
        if isinstance(device, (list, tuple)):
            self.devices = [int(d) for d in device]
            self.device = self.devices[0]
            arr = (C.c_int * len(self.devices))(*self.devices)
            nat.check(self._lib.mcp_index_create_multi(C.byref(self._h), arr, len(self.devices), self.dim, self.metric))
        else:
            self.devices = [int(device)]
            self.device = int(device)
            nat.check(self._lib.mcp_index_create(C.byref(self._h), self.device, self.dim, self.metric))
        self._keep = []  # adopted tensors must outlive the handle

    # ---- lifetime -----------------------------------------------------------------------
    def close(self) -> None:
        if self._h:
            self._lib.mcp_index_destroy(self._h)
            self._h = C.c_void_p()
            self._keep.clear()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def n_rows(self) -> int:
        return int(self._lib.mcp_index_rows(self._h))

    # ---- table ---------------------------------------------------------------------------
    def load(self, rows: np.ndarray, id_base: int = 0) -> "ServiceIndex":
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        if rows.ndim != 2 or rows.shape[1] != self.dim:
            raise ValueError(f"rows must be [N, {self.dim}] float32, got {rows.shape}")
        nat.check(self._lib.mcp_index_load(self._h, _np_ptr(rows), rows.shape[0], int(id_base)))
        self._keep.clear()
        return self

    def adopt(self, rows, id_base: int = 0) -> "ServiceIndex":
        """Use a CUDA float32 tensor [N, dim] (row stride >= dim) in place; the caller must not modify it."""
        import torch

        if not (rows.is_cuda and rows.dtype == torch.float32 and rows.dim() == 2 and rows.stride(1) == 1):
            raise ValueError("adopt needs a CUDA float32 [N, dim] tensor with unit column stride")
        if rows.shape[1] != self.dim or rows.device.index != self.device:
            raise ValueError("adopt: wrong dim or device")
        nat.check(self._lib.mcp_index_adopt_device(self._h, C.c_void_p(rows.data_ptr()), rows.shape[0],
                                                   rows.stride(0), int(id_base), _stream_ptr()))
        self._keep = [rows]
        return self

    def update_rows(self, first_row: int, rows: Optional[np.ndarray] = None, n_rows: Optional[int] = None) -> "ServiceIndex":
        """Replace rows [first_row, first_row + len(rows)) of a loaded table, or (rows=None) tell the index that the owner
        of an adopted table changed n_rows rows in place; norms are recomputed, nothing else is reloaded."""
        if rows is not None:
            rows = np.ascontiguousarray(rows, dtype=np.float32)
            if rows.ndim != 2 or rows.shape[1] != self.dim:
                raise ValueError(f"rows must be [n, {self.dim}] float32")
            nat.check(self._lib.mcp_index_update_rows(self._h, int(first_row), _np_ptr(rows), rows.shape[0]))
        else:
            nat.check(self._lib.mcp_index_update_rows(self._h, int(first_row), None, int(n_rows)))
        return self

    def upsert_rows(self, row_ids, rows: np.ndarray) -> "ServiceIndex":
        """Replace the scattered rows ``row_ids`` (row numbers) of a loaded table without stopping queries
        (``mcp_index_upsert_rows``): returns once the update is enqueued behind the searches in flight."""
        ids = np.ascontiguousarray(row_ids, dtype=np.int64)
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        if rows.ndim != 2 or rows.shape != (ids.shape[0], self.dim):
            raise ValueError(f"rows must be [{ids.shape[0]}, {self.dim}] float32")
        nat.check(self._lib.mcp_index_upsert_rows(self._h, _np_ptr(ids), _np_ptr(rows), ids.shape[0]))
        return self

    # ---- flat-file table format ------------------------------------------------
    # <path>: 64-byte header (magic "MCPTBL01", dim, n_rows, id_base, metric as little-endian int64) + N x dim fp32 rows,
    # row-major -- mmap-able, so a registry loads at file-system speed instead of through Postgres rows (control_plane.py:54).
    # <path>.names: one service name per line (optional sidecar).
    @staticmethod
    def save_table(path, rows: np.ndarray, names=None, id_base: int = 0, metric: int = 0) -> None:
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        hdr = np.zeros(8, np.int64)
        hdr[0] = int.from_bytes(b"MCPTBL01", "little")
        hdr[1:5] = (rows.shape[1], rows.shape[0], id_base, metric)
        with open(path, "wb") as f:
            f.write(hdr.tobytes())
            f.write(rows.tobytes())
        if names is not None:
            with open(str(path) + ".names", "w", encoding="utf-8") as f:
                f.write("\n".join(names))

    @staticmethod
    def open_table(path):
        """-> (rows memmap [N, dim] float32, names or None, id_base, metric) without reading the rows."""
        hdr = np.fromfile(path, dtype=np.int64, count=8)
        if hdr.size != 8 or hdr[0] != int.from_bytes(b"MCPTBL01", "little"):
            raise ValueError(f"{path}: not an MCPTBL01 table file")
        dim, n, id_base, metric = (int(x) for x in hdr[1:5])
        rows = np.memmap(path, dtype=np.float32, mode="r", offset=64, shape=(n, dim))
        names = None
        try:
            with open(str(path) + ".names", encoding="utf-8") as f:
                names = f.read().split("\n")
        except FileNotFoundError:
            pass
        return rows, names, id_base, metric

    def load_file(self, path, first_row: int = 0, n_rows: Optional[int] = None) -> list:
        """Load rows ``[first_row, first_row + n_rows)`` of a table written by ``save_table`` (default: all of it) through the
        library's pinned chunk pipeline (``mcp_index_load_file``); service ids are the file's id_base + row number in the
        file.  Returns the names sidecar (or None), sliced to the same rows."""
        nat.check(self._lib.mcp_index_load_file(self._h, str(path).encode(), int(first_row), -1 if n_rows is None else int(n_rows)))
        self._keep.clear()
        names = None
        try:
            with open(str(path) + ".names", encoding="utf-8") as f:
                names = f.read().split("\n")[first_row:first_row + self.n_rows]
        except FileNotFoundError:
            pass
        return names

    def set_penalty(self, penalty) -> "ServiceIndex":
        if penalty is None:
            nat.check(self._lib.mcp_index_set_penalty(self._h, None))
        elif isinstance(penalty, np.ndarray):
            p = np.ascontiguousarray(penalty, dtype=np.float32)
            if p.shape != (self.n_rows,):
                raise ValueError(f"penalty must be [{self.n_rows}]")
            nat.check(self._lib.mcp_index_set_penalty(self._h, _np_ptr(p)))
        else:
            import torch

            if not (penalty.is_cuda and penalty.dtype == torch.float32 and penalty.is_contiguous()
                    and tuple(penalty.shape) == (self.n_rows,)):
                raise ValueError("penalty tensor must be contiguous CUDA float32 [N]")
            nat.check(self._lib.mcp_index_set_penalty_device(self._h, C.c_void_p(penalty.data_ptr()), _stream_ptr()))
        return self

    # ---- search --------------------------------------------------------------------------
    def search(self, queries: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        """Host buffers in and out: (ids int32 [Q,k], scores fp32 [Q,k], scores fp64 [Q,k])."""
        q = np.ascontiguousarray(np.atleast_2d(queries), dtype=np.float32)
        if q.shape[1] != self.dim:
            raise ValueError(f"queries must be [Q, {self.dim}]")
        nq = q.shape[0]
        ids = np.empty((nq, k), np.int32)
        sc = np.empty((nq, k), np.float32)
        sc64 = np.empty((nq, k), np.float64)
        nat.check(self._lib.mcp_index_search(self._h, _np_ptr(q), nq, int(k), _np_ptr(ids), _np_ptr(sc), _np_ptr(sc64)))
        return ids, sc, sc64

    def search_device(self, queries, k: int, out: Optional[tuple] = None):
        """CUDA tensors in and out, enqueued on the current stream; call ``sync_status`` to learn
        whether every answer was certified."""
        import torch

        if not (queries.is_cuda and queries.dtype == torch.float32 and queries.is_contiguous()
                and queries.dim() == 2 and queries.shape[1] == self.dim and queries.device.index == self.device):
            raise ValueError("queries must be contiguous CUDA float32 [Q, dim] on the index's device")
        nq = queries.shape[0]
        if out is None:
            dev = queries.device
            out = (torch.empty((nq, k), dtype=torch.int32, device=dev),
                   torch.empty((nq, k), dtype=torch.float32, device=dev),
                   torch.empty((nq, k), dtype=torch.float64, device=dev))
        ids, sc, sc64 = out
        nat.check(self._lib.mcp_index_search_device(
            self._h, C.c_void_p(queries.data_ptr()), nq, int(k), C.c_void_p(ids.data_ptr()),
            C.c_void_p(sc.data_ptr()), C.c_void_p(sc64.data_ptr()) if sc64 is not None else None, _stream_ptr()))
        return ids, sc, sc64

    def repair_device(self, queries, k: int, out: tuple):
        """After ``sync_status`` reported unresolved queries for the ``search_device`` call that produced ``out``: re-answers
        exactly those (dense fp64 scan) in place.  Blocks."""
        ids, sc, sc64 = out
        nat.check(self._lib.mcp_index_repair_device(
            self._h, C.c_void_p(queries.data_ptr()), queries.shape[0], int(k), C.c_void_p(ids.data_ptr()),
            C.c_void_p(sc.data_ptr()), C.c_void_p(sc64.data_ptr()) if sc64 is not None else None, _stream_ptr()))
        return out

    def search_exact_noalloc_device(self, queries, k: int, out: tuple):
        """fp64 full scan through ``mcp_index_search_exact_noalloc_device``: the library allocates nothing, so the
        caller hands in the outputs too -- ``out = (ids int32 [Q,k], scores fp32 [Q,k], scores fp64 [Q,k])`` CUDA
        tensors made ahead of time.  For request paths that must not hit cudaMalloc (it would stall a scan
        running on another stream, or fail when HBM is fully budgeted).  At most 1,048,576 rows and 64 queries;
        all calls on one device share one fixed scratch array, so keep them on one stream.  The first call on a
        device does a one-time address look-up: make it before starting any stream capture."""
        import torch

        ids, sc, sc64 = out
        if not (queries.is_cuda and queries.dtype == torch.float32 and queries.is_contiguous() and queries.dim() == 2
                and queries.shape[1] == self.dim and queries.device.index == self.device):
            raise ValueError("queries must be contiguous CUDA float32 [Q, dim] on the index's device")
        nq = queries.shape[0]
        for t, dt in ((ids, torch.int32), (sc, torch.float32), (sc64, torch.float64)):
            if not (t.is_cuda and t.dtype == dt and t.is_contiguous() and tuple(t.shape) == (nq, k) and t.device == queries.device):
                raise ValueError("out = (ids int32, scores float32, scores float64), contiguous CUDA [Q, k] tensors on the index's device")
        nat.check(self._lib.mcp_index_search_exact_noalloc_device(
            self._h, C.c_void_p(queries.data_ptr()), nq, int(k), C.c_void_p(ids.data_ptr()),
            C.c_void_p(sc.data_ptr()), C.c_void_p(sc64.data_ptr()), _stream_ptr()))
        return ids, sc, sc64

    def coarse_scores_device(self, queries, path="gemm_tf32"):
        """[Q, N] fp32 coarse scores exactly as one coarse kernel computes them (validation aid)."""
        import torch

        out = torch.empty((queries.shape[0], self.n_rows), dtype=torch.float32, device=queries.device)
        nat.check(self._lib.mcp_index_coarse_scores_device(
            self._h, _PATHS[path], C.c_void_p(queries.data_ptr()), queries.shape[0], C.c_void_p(out.data_ptr()),
            _stream_ptr()))
        return out

    def sync_status(self) -> Tuple[int, int]:
        a, b = C.c_int32(), C.c_int32()
        nat.check(self._lib.mcp_index_sync_status(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # ---- knobs ---------------------------------------------------------------------------
    def set_option(self, name: str, value) -> "ServiceIndex":
        if name == "path" and isinstance(value, str):
            value = _PATHS[value]
        nat.check(self._lib.mcp_index_set_option(self._h, name.encode(), int(value)))
        return self

    def stats(self) -> dict:
        s = nat.McpStats()
        nat.check(self._lib.mcp_index_get_stats(self._h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in s._fields_}


def merge_topk_device(scores64, ids, out: Optional[tuple] = None):
    """[S,Q,k] per-shard sorted lists (CUDA fp64 / int32) -> global (ids, scores32, scores64)."""
    import torch

    s, q, k = ids.shape
    if not (scores64.is_cuda and scores64.dtype == torch.float64 and scores64.is_contiguous()
            and ids.dtype == torch.int32 and ids.is_contiguous() and tuple(scores64.shape) == (s, q, k)):
        raise ValueError("merge needs contiguous CUDA [S,Q,k] fp64 scores and int32 ids")
    if out is None:
        dev = ids.device
        out = (torch.empty((q, k), dtype=torch.int32, device=dev),
               torch.empty((q, k), dtype=torch.float32, device=dev),
               torch.empty((q, k), dtype=torch.float64, device=dev))
    oi, os32, os64 = out
    nat.check(nat.lib().mcp_merge_topk_device(C.c_void_p(scores64.data_ptr()), C.c_void_p(ids.data_ptr()), s, q, k,
                                              C.c_void_p(oi.data_ptr()), C.c_void_p(os32.data_ptr()),
                                              C.c_void_p(os64.data_ptr()), _stream_ptr()))
    return oi, os32, os64


def packed_block_bytes(n_queries: int, k: int) -> int:
    """Bytes of one shard's block in the packed gather buffer (``mcp_merge_topk_packed_device``)."""
    return int(nat.lib().mcp_merge_packed_block_bytes(int(n_queries), int(k)))


def packed_views(block, n_queries: int, k: int):
    """(ids int32 [Q,k], scores fp64 [Q,k]) views into one shard's block (a CUDA uint8 tensor of ``packed_block_bytes``)."""
    import torch

    qk = n_queries * k
    sc64 = block[: qk * 8].view(torch.float64).view(n_queries, k)
    ids = block[qk * 8: qk * 12].view(torch.int32).view(n_queries, k)
    return ids, sc64


def merge_topk_packed_device(packed, n_shards: int, n_queries: int, k: int, out: Optional[tuple] = None):
    """[S * block] gathered uint8 buffer -> global (ids, scores32, scores64)."""
    import torch

    if not (packed.is_cuda and packed.dtype == torch.uint8 and packed.is_contiguous()
            and packed.numel() == n_shards * packed_block_bytes(n_queries, k)):
        raise ValueError("packed merge needs a contiguous CUDA uint8 buffer of n_shards blocks")
    if out is None:
        dev = packed.device
        out = (torch.empty((n_queries, k), dtype=torch.int32, device=dev),
               torch.empty((n_queries, k), dtype=torch.float32, device=dev),
               torch.empty((n_queries, k), dtype=torch.float64, device=dev))
    oi, os32, os64 = out
    nat.check(nat.lib().mcp_merge_topk_packed_device(C.c_void_p(packed.data_ptr()), n_shards, n_queries, k,
                                                     C.c_void_p(oi.data_ptr()), C.c_void_p(os32.data_ptr()),
                                                     C.c_void_p(os64.data_ptr()), _stream_ptr()))
    return oi, os32, os64

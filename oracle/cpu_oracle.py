"""ctypes face of ``oracle.c`` plus the two Python references that pin / time it.

* ``topk`` / ``scores`` / ``merge`` / ``dag_levels``  -> the C restatement (fixed-order fp64).
* ``topk_python``    -> the same arithmetic in pure Python loops; only for tiny cases, used to
                        check that the C file means what its comments say.
* ``nx_levels``      -> the live NetworkX call the reference makes (control_plane.py:94-104).
* ``numpy_topk_baseline`` -> BASELINE.md section 3's timed CPU path (OpenBLAS sgemm + argpartition);
                        a *baseline*, not the correctness oracle (fp32, summation order is BLAS's).
* ``telemetry_penalty`` -> the one fp32 numpy expression both the oracle and the GPU path are fed.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from pathlib import Path

import numpy as np

COSINE, L2 = 0, 1
_HERE = Path(__file__).resolve().parent
_SO = _HERE / "_build" / "liboracle.so"
_lib = None


def build(force: bool = False) -> Path:
    """Compile oracle.c with gcc (idempotent). Building the checker is not using it."""
    src = _HERE / "oracle.c"
    if force or not _SO.exists() or _SO.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "-s", "-B"], check=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(str(_SO))
        c = ctypes
        fp, dp, ip = c.POINTER(c.c_float), c.POINTER(c.c_double), c.POINTER(c.c_int32)
        u8, u32 = c.POINTER(c.c_uint8), c.POINTER(c.c_uint32)
        lib.oracle_scores.argtypes = [fp, c.c_int64, c.c_int, c.c_int64, fp, c.c_int, fp, dp]
        lib.oracle_scores.restype = None
        lib.oracle_pgvector_f32_scores.argtypes = [fp, c.c_int64, c.c_int, c.c_int64, fp, c.c_int, fp, dp]
        lib.oracle_pgvector_f32_scores.restype = None
        lib.oracle_topk.argtypes = [fp, c.c_int64, c.c_int, c.c_int64, fp, c.c_int, c.c_int, c.c_int, fp,
                                    c.c_int64, ip, dp, fp]
        lib.oracle_topk.restype = None
        lib.oracle_merge.argtypes = [dp, ip, c.c_int, c.c_int, c.c_int, ip, dp]
        lib.oracle_merge.restype = None
        lib.oracle_dag_levels.argtypes = [u8, u32, u8, c.c_int64, u8, u8, u8, u8]
        lib.oracle_dag_levels.restype = None
        _lib = lib
    return _lib


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t)) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _chunked(fn, n, threads=None):
    """Run fn(lo, hi) over [0, n) split across host threads (the C file itself is scalar)."""
    from concurrent.futures import ThreadPoolExecutor

    t = max(1, min(threads or cpu_count(), n))
    if t == 1:
        fn(0, n)
        return
    bounds = np.linspace(0, n, t + 1).astype(np.int64)
    with ThreadPoolExecutor(t) as ex:
        list(ex.map(lambda i: fn(int(bounds[i]), int(bounds[i + 1])), range(t)))


def scores(rows, query, metric=COSINE, penalty=None) -> np.ndarray:
    rows, query = _f32(rows), _f32(query)
    n, d = rows.shape
    pen = _f32(penalty) if penalty is not None else None
    out = np.empty(n, np.float64)
    _load().oracle_scores(_p(rows, ctypes.c_float), n, d, d, _p(query, ctypes.c_float), metric,
                          _p(pen, ctypes.c_float), _p(out, ctypes.c_double))
    return out


def pgvector_f32_scores(rows, query, metric=COSINE, penalty=None) -> np.ndarray:
    """Scores with pgvector's own arithmetic (float accumulators in index order, see oracle.c): the "exact fp32 brute-force
    reference" north_star measures the 1e-5 score tolerance against.  Not the oracle."""
    rows, query = _f32(rows), _f32(query)
    n, d = rows.shape
    pen = _f32(penalty) if penalty is not None else None
    out = np.empty(n, np.float64)
    _load().oracle_pgvector_f32_scores(_p(rows, ctypes.c_float), n, d, d, _p(query, ctypes.c_float), metric,
                                       _p(pen, ctypes.c_float), _p(out, ctypes.c_double))
    return out


def agrees_with_fp32_reference(ids, scores64, ref_scores, k, rel=1e-5):
    """True when a top-k answer (ids [k], fp64 scores [k]) is the answer of the fp32 brute-force reference up to its own
    rounding: every returned score within ``rel`` of the reference's score of the same row, and every row the reference
    ranks in its top k that is missing from ``ids`` (or the other way round) within ``rel`` of the reference's k-th score."""
    ids = np.asarray(ids)
    ok = ids >= 0
    if not np.allclose(scores64[ok], ref_scores[ids[ok]], rtol=rel, atol=0):
        return False
    order = np.lexsort((np.arange(ref_scores.size), ref_scores))[:k]
    kth = ref_scores[order[min(k, order.size) - 1]]
    diff = np.setxor1d(ids[ok], order)
    return bool(np.all(np.abs(ref_scores[diff] - kth) <= rel * max(abs(kth), 1e-30)))


def topk(rows, queries, k, metric=COSINE, penalty=None, id_base=0, threads=None):
    """-> (ids int32 [Q,k], scores64 [Q,k], scores32 [Q,k]); ranking key (score_f64, id)."""
    rows, queries = _f32(rows), np.atleast_2d(_f32(queries))
    n, d = rows.shape
    nq = queries.shape[0]
    pen = _f32(penalty) if penalty is not None else None
    ids = np.empty((nq, k), np.int32)
    s64 = np.empty((nq, k), np.float64)
    s32 = np.empty((nq, k), np.float32)
    lib = _load()

    def run(lo, hi):  # ctypes drops the GIL, so query chunks run on separate cores
        lib.oracle_topk(_p(rows, ctypes.c_float), n, d, d, _p(queries[lo:hi], ctypes.c_float), hi - lo, k,
                        metric, _p(pen, ctypes.c_float), id_base, _p(ids[lo:hi], ctypes.c_int32),
                        _p(s64[lo:hi], ctypes.c_double), _p(s32[lo:hi], ctypes.c_float))

    _chunked(run, nq, threads)
    return ids, s64, s32


def merge(scores64, ids):
    """[S,Q,k] sorted per-shard lists -> global (ids [Q,k], scores64 [Q,k])."""
    scores64 = np.ascontiguousarray(scores64, np.float64)
    ids = np.ascontiguousarray(ids, np.int32)
    s, q, k = ids.shape
    oi = np.empty((q, k), np.int32)
    os_ = np.empty((q, k), np.float64)
    _load().oracle_merge(_p(scores64, ctypes.c_double), _p(ids, ctypes.c_int32), s, q, k,
                         _p(oi, ctypes.c_int32), _p(os_, ctypes.c_double))
    return oi, os_


def dag_levels(n_nodes, edge_off, edges, threads=None):
    """Encoded batch -> (level [B,64] u8, order [B,64] u8, n_levels [B] u8, status [B] u8)."""
    n_nodes = np.ascontiguousarray(n_nodes, np.uint8)
    edge_off = np.ascontiguousarray(edge_off, np.uint32)
    edges = np.ascontiguousarray(edges, np.uint8).reshape(-1)
    b = n_nodes.shape[0]
    level = np.empty((b, 64), np.uint8)
    order = np.empty((b, 64), np.uint8)
    nlev = np.empty(b, np.uint8)
    status = np.empty(b, np.uint8)
    if edges.size == 0:
        edges = np.zeros(2, np.uint8)
    lib = _load()

    def run(lo, hi):  # edge_off is absolute, so a chunk only needs shifted per-DAG pointers
        lib.oracle_dag_levels(_p(n_nodes[lo:hi], ctypes.c_uint8), _p(edge_off[lo:hi + 1], ctypes.c_uint32),
                              _p(edges, ctypes.c_uint8), hi - lo, _p(level[lo:hi], ctypes.c_uint8),
                              _p(order[lo:hi], ctypes.c_uint8), _p(nlev[lo:hi], ctypes.c_uint8),
                              _p(status[lo:hi], ctypes.c_uint8))

    _chunked(run, b, threads)
    return level, order, nlev, status


# ---------------------------------------------------------------------------------------------
# pure-Python restatement (tiny cases only)
# ---------------------------------------------------------------------------------------------
def topk_python(rows, queries, k, metric=COSINE, penalty=None):
    rows = np.asarray(rows, np.float32)
    queries = np.atleast_2d(np.asarray(queries, np.float32))
    out_ids, out_s = [], []
    for q in queries:
        keyed = []
        nq2 = 0.0
        for x in q:
            nq2 = nq2 + float(x) * float(x)
        for r, e in enumerate(rows):
            if metric == COSINE:
                dot = ne2 = 0.0
                for a, b in zip(q, e):
                    a, b = float(a), float(b)
                    dot = dot + a * b
                    ne2 = ne2 + b * b
                den = math.sqrt(nq2 * ne2)
                sim = dot / den if den != 0.0 else float("nan")
                if sim > 1.0:
                    sim = 1.0
                elif sim < -1.0:
                    sim = -1.0
                d = 1.0 - sim
            else:
                s = 0.0
                for a, b in zip(q, e):
                    t = float(a) - float(b)
                    s = s + t * t
                d = math.sqrt(s)
            if penalty is not None:
                d = d + float(np.float32(penalty[r]))
            keyed.append(((1, 0.0) if math.isnan(d) else (0, d), r, d))
        keyed.sort(key=lambda t: (t[0], t[1]))
        keyed = keyed[:k]
        ids = [t[1] for t in keyed] + [-1] * (k - len(keyed))
        sc = [t[2] for t in keyed] + [float("inf")] * (k - len(keyed))
        out_ids.append(ids)
        out_s.append(sc)
    return np.array(out_ids, np.int32), np.array(out_s, np.float64)


# ---------------------------------------------------------------------------------------------
# the live reference for level scheduling
# ---------------------------------------------------------------------------------------------
def nx_levels(graph: dict):
    """Build the DiGraph exactly as control_plane.py:94-100 does and run the call behind :104.

    -> (levels {name: generation}, order [names as topological_sort yields them], status)
    status 1 means NetworkXUnfeasible was raised; levels/order then hold the acyclic prefix.
    """
    import networkx as nx

    G = nx.DiGraph()
    for node in graph["nodes"]:
        G.add_node(node["name"], **node)
    for edge in graph["edges"]:
        G.add_edge(edge["from"], edge["to"], fallback=edge.get("fallback"))
    levels, order, status = {}, [], 0
    try:
        for i, gen in enumerate(nx.topological_generations(G)):
            for n in gen:
                levels[n] = i
                order.append(n)
    except nx.NetworkXUnfeasible:
        status = 1
    return levels, order, status


# ---------------------------------------------------------------------------------------------
# telemetry fold + timed CPU baseline
# ---------------------------------------------------------------------------------------------
def telemetry_penalty(latency_ms, error_rate, cost_profile, w_lat=0.05, w_err=0.5, w_cost=1.0,
                      lat_ref=1000.0) -> np.ndarray:
    """one fp32 penalty per service, computed once on the host."""
    lat = np.asarray(latency_ms, np.float32)
    err = np.asarray(error_rate, np.float32)
    cost = np.asarray(cost_profile, np.float32)
    one = np.float32(1.0)
    pen = (np.float32(w_lat) * np.minimum(lat / np.float32(lat_ref), one)
           + np.float32(w_err) * err + np.float32(w_cost) * cost)
    return pen.astype(np.float32)


def numpy_topk_baseline(rows, inv_norm_rows, queries, k, metric=COSINE, penalty=None, row_norm2=None):
    """S = Q @ E.T (OpenBLAS sgemm) -> distance -> + penalty -> argpartition."""
    q = np.asarray(queries, np.float32)
    s = q @ rows.T
    if metric == COSINE:
        inv_q = (1.0 / np.sqrt(np.einsum("ij,ij->i", q, q))).astype(np.float32)
        f = 1.0 - s * inv_q[:, None] * inv_norm_rows[None, :]
    else:
        nq = np.einsum("ij,ij->i", q, q).astype(np.float32)
        f = np.sqrt(np.maximum(nq[:, None] + row_norm2[None, :] - 2.0 * s, 0.0))
    if penalty is not None:
        f = f + penalty[None, :]
    kk = min(k, f.shape[1])
    part = np.argpartition(f, kk - 1, axis=1)[:, :kk]
    ps = np.take_along_axis(f, part, axis=1)
    order = np.lexsort((part, ps), axis=1)
    ids = np.take_along_axis(part, order, axis=1).astype(np.int32)
    return ids, np.take_along_axis(ps, order, axis=1)


def numpy_topk_blocked(rows, inv_norm_rows, queries, k, metric=COSINE, penalty=None, row_norm2=None, block=65536):
    """``numpy_topk_baseline`` over row blocks of ``block`` rows, the running top-k carried from block to block (the
    [Q, N] score matrix of a 1024-query batch over a million rows would not fit the caches, or comfortably in RAM).
    Same arithmetic per element; the timed CPU arm of ``bench.py``."""
    n = rows.shape[0]
    if n <= block:
        return numpy_topk_baseline(rows, inv_norm_rows, queries, k, metric, penalty, row_norm2)
    best_i = best_s = None
    for lo in range(0, n, block):
        hi = min(n, lo + block)
        ids, sc = numpy_topk_baseline(rows[lo:hi], inv_norm_rows[lo:hi], queries, k, metric,
                                      None if penalty is None else penalty[lo:hi], None if row_norm2 is None else row_norm2[lo:hi])
        ids = ids + np.int32(lo)
        if best_i is None:
            best_i, best_s = ids, sc
        else:
            ci, cs = np.concatenate([best_i, ids], axis=1), np.concatenate([best_s, sc], axis=1)
            order = np.lexsort((ci, cs), axis=1)[:, :k]
            best_i, best_s = np.take_along_axis(ci, order, axis=1), np.take_along_axis(cs, order, axis=1)
    return best_i, best_s


def cpu_count() -> int:
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)

"""Plan-DAG encoding and batched level scheduling (host side of ``mcp_dag_levels``).

``encode_graphs`` turns plan dicts of the shape ``Orchestrator.execute`` receives
(control_plane.py:93-100) into the integer arrays the kernel reads, numbering nodes the way
``nx.DiGraph`` would hold them:

  * ``for node in graph["nodes"]: G.add_node(node["name"], **node)``  (:96-97) -- declared nodes
    in order, a repeated name keeps its first position;
  * ``G.add_edge(edge["from"], edge["to"], ...)``  (:98-100) -- an endpoint that was never
    declared is appended when first seen (``from`` before ``to``); repeated edges stay in the
    edge list (the kernel collapses them like a DiGraph does).

``levels`` then gives, per graph, what ``enumerate(nx.topological_generations(G))`` gives and the
order ``nx.topological_sort(G)`` (:104) walks.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Iterable, List, Sequence, Tuple

import numpy as np

from . import _native as nat

MAX_NODES = nat.MAX_DAG_NODES            # the batched warp-per-DAG kernel
MAX_WIDE_NODES = nat.MAX_WIDE_DAG_NODES  # single oversized plans: the CTA-per-DAG kernel


@dataclass
class EncodedDags:
    n_nodes: np.ndarray          # uint8  [B]
    edge_off: np.ndarray         # uint32 [B+1]
    edges: np.ndarray            # uint8  [E, 2]  (src, dst)
    names: List[List[str]]       # node index -> name, per graph (empty when synthetic)

    def __len__(self) -> int:
        return int(self.n_nodes.shape[0])


@dataclass
class DagLevels:
    level: np.ndarray            # uint8 [B, 64]; 0xFF = no node / never emitted
    order: np.ndarray            # uint8 [B, 64]; 0xFF past the end
    n_levels: np.ndarray         # uint8 [B]
    status: np.ndarray           # uint8 [B]; 1 = cycle (NetworkXUnfeasible)


def encode_graphs(graphs: Iterable[dict]) -> EncodedDags:
    n_nodes, offs, edges, names = [], [0], [], []
    for gi, g in enumerate(graphs):
        index: Dict[str, int] = {}
        for node in g["nodes"]:
            index.setdefault(node["name"], len(index))
        for e in g["edges"]:
            u = index.setdefault(e["from"], len(index))
            v = index.setdefault(e["to"], len(index))
            edges.append((u, v))
        if len(index) > MAX_WIDE_NODES:
            raise ValueError(f"graph {gi} has {len(index)} nodes; the level kernels handle at most {MAX_WIDE_NODES}")
        n_nodes.append(len(index))
        offs.append(len(edges))
        names.append(list(index))
    return EncodedDags(
        np.asarray(n_nodes, np.uint8),
        np.asarray(offs, np.uint32),
        np.asarray(edges, np.uint8).reshape(-1, 2),
        names,
    )


def _levels_call(fn, pitch, nn, off, ed, device, with_order):
    b = nn.shape[0]
    level = np.empty((b, pitch), np.uint8)
    order = np.empty((b, pitch), np.uint8) if with_order else None
    nlev = np.empty(b, np.uint8)
    status = np.empty(b, np.uint8)
    nat.check(fn(int(device), C.c_void_p(nn.ctypes.data), C.c_void_p(off.ctypes.data),
                 C.c_void_p(ed.ctypes.data) if ed.size else None, b, C.c_void_p(level.ctypes.data),
                 C.c_void_p(order.ctypes.data) if with_order else None, C.c_void_p(nlev.ctypes.data),
                 C.c_void_p(status.ctypes.data)))
    return level, order, nlev, status


def levels(enc: EncodedDags, device: int = 0, with_order: bool = True) -> DagLevels:
    """Host arrays in, host arrays out, through ``mcp_dag_levels`` (copies inside).  Graphs of more than 64 nodes (up to
    255) go through ``mcp_dag_levels_wide``; the result arrays are then 256 wide instead of 64."""
    b = len(enc)
    nn = np.ascontiguousarray(enc.n_nodes, np.uint8)
    off = np.ascontiguousarray(enc.edge_off, np.uint32)
    ed = np.ascontiguousarray(enc.edges, np.uint8)
    lib = nat.lib()
    wide = nn > MAX_NODES
    if not b:
        return DagLevels(np.empty((0, 64), np.uint8), np.empty((0, 64 if with_order else 0), np.uint8), np.empty(0, np.uint8), np.empty(0, np.uint8))
    if not wide.any():
        level, order, nlev, status = _levels_call(lib.mcp_dag_levels, 64, nn, off, ed, device, with_order)
        return DagLevels(level, order if with_order else np.zeros((b, 0), np.uint8), nlev, status)
    # mixed batch: the oversized graphs are pulled out into a batch of their own
    pitch = 256
    level = np.full((b, pitch), 0xFF, np.uint8)
    order = np.full((b, pitch), 0xFF, np.uint8) if with_order else np.zeros((b, 0), np.uint8)
    nlev = np.zeros(b, np.uint8)
    status = np.zeros(b, np.uint8)
    off64 = off.astype(np.int64)
    for sel, fn, p in ((np.flatnonzero(~wide), lib.mcp_dag_levels, 64), (np.flatnonzero(wide), lib.mcp_dag_levels_wide, 256)):
        if not sel.size:
            continue
        cnt = off64[sel + 1] - off64[sel]
        sub_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint32)
        sub_ed = np.concatenate([ed[off64[i]:off64[i + 1]] for i in sel]) if cnt.sum() else np.zeros((0, 2), np.uint8)
        lv, od, nl, st = _levels_call(fn, p, np.ascontiguousarray(nn[sel]), sub_off, np.ascontiguousarray(sub_ed), device, with_order)
        level[sel, :p] = lv
        if with_order:
            order[sel, :p] = od
        nlev[sel], status[sel] = nl, st
    return DagLevels(level, order, nlev, status)


def launch_counts() -> Dict[str, int]:
    """Kernel launches so far in this process behind the <= 64-node entry points (``mcp_dag_launch_count``)."""
    lib = nat.lib()
    return {name: int(lib.mcp_dag_launch_count(i)) for i, name in enumerate(("warp", "thread", "handed_back"))}


def levels_device(n_nodes, edge_off, edges, out=None, with_order: bool = True):
    """CUDA uint8 / int32-viewed-as-uint32 tensors in and out on the current stream (no copies)."""
    import torch

    b = n_nodes.shape[0]
    dev = n_nodes.device
    if out is None:
        out = (torch.empty((b, 64), dtype=torch.uint8, device=dev),
               torch.empty((b, 64), dtype=torch.uint8, device=dev) if with_order else None,
               torch.empty(b, dtype=torch.uint8, device=dev),
               torch.empty(b, dtype=torch.uint8, device=dev))
    level, order, nlev, status = out
    nat.check(nat.lib().mcp_dag_levels_device(
        C.c_void_p(n_nodes.data_ptr()), C.c_void_p(edge_off.data_ptr()), C.c_void_p(edges.data_ptr()), b,
        C.c_void_p(level.data_ptr()), C.c_void_p(order.data_ptr()) if order is not None else None,
        C.c_void_p(nlev.data_ptr()), C.c_void_p(status.data_ptr()),
        C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def decode(enc: EncodedDags, res: DagLevels, i: int) -> Tuple[Dict[str, int], List[str], int]:
    """Graph i -> ({name: generation}, [names in topological_sort order], status)."""
    names = enc.names[i]
    lv = {names[v]: int(res.level[i, v]) for v in range(len(names)) if res.level[i, v] != 0xFF}
    order = [names[v] for v in res.order[i] if v != 0xFF] if res.order.shape[1] else []
    return lv, order, int(res.status[i])


def generations(enc: EncodedDags, res: DagLevels, i: int) -> List[List[str]]:
    """The list ``list(nx.topological_generations(G))`` would give (acyclic prefix on a cycle)."""
    lv, order, _ = decode(enc, res, i)
    gens: List[List[str]] = [[] for _ in range(int(res.n_levels[i]))]
    for name in order:
        gens[lv[name]].append(name)
    return gens


def synthetic_dags(n_dags: int, seed: int = 4004, chunk: int = 10000) -> EncodedDags:
    """n ~ U{1..64}; a random topological permutation; every forward
    pair is an edge with probability p ~ U(0.02, 0.25) per DAG; edges shuffled; 1 % of DAGs get
    one back edge (a cycle)."""
    rng = np.random.default_rng(seed)
    nn_all, off_all, ed_all = [], [np.zeros(1, np.int64)], []
    total = 0
    iu, ju = np.triu_indices(64, 1)
    for c0 in range(0, n_dags, chunk):
        b = min(chunk, n_dags - c0)
        n = rng.integers(1, 65, b)
        p = rng.uniform(0.02, 0.25, b)
        perm = np.argsort(rng.random((b, 64)), axis=1).astype(np.uint8)   # position -> label (labels >= n filtered)
        keep = rng.random((b, iu.size), dtype=np.float32) < p[:, None].astype(np.float32)
        lab_ok = perm < n[:, None]                                        # labels >= n are not nodes of this DAG
        valid_pair = lab_ok[:, iu] & lab_ok[:, ju] & keep
        src = perm[:, iu][valid_pair]
        dst = perm[:, ju][valid_pair]
        dag = np.broadcast_to(np.arange(b)[:, None], valid_pair.shape)[valid_pair]
        # 1 % back edges: reverse one existing edge's direction as an extra edge
        cyc = rng.random(b) < 0.01
        counts = np.bincount(dag, minlength=b)
        starts = np.concatenate([[0], np.cumsum(counts)[:-1]])
        pick = np.where(cyc & (counts > 0))[0]
        extra_src = dst[starts[pick]]
        extra_dst = src[starts[pick]]
        dag = np.concatenate([dag, pick])
        src = np.concatenate([src, extra_src])
        dst = np.concatenate([dst, extra_dst])
        # shuffle edges within each DAG
        key = dag.astype(np.float64) + rng.random(dag.size) * 0.999
        o = np.argsort(key, kind="stable")
        dag, src, dst = dag[o], src[o], dst[o]
        counts = np.bincount(dag, minlength=b)
        nn_all.append(n.astype(np.uint8))
        off_all.append(total + np.cumsum(counts))
        total += int(counts.sum())
        ed_all.append(np.stack([src, dst], axis=1).astype(np.uint8))
    return EncodedDags(np.concatenate(nn_all), np.concatenate(off_all).astype(np.uint32),
                       np.concatenate(ed_all) if ed_all else np.zeros((0, 2), np.uint8), [])


def to_graph_dict(enc: EncodedDags, i: int) -> dict:
    """Encoded graph i back to the reference's plan-dict shape (for checking against NetworkX)."""
    n = int(enc.n_nodes[i])
    names = enc.names[i] if enc.names else [f"n{v}" for v in range(n)]
    e0, e1 = int(enc.edge_off[i]), int(enc.edge_off[i + 1])
    return {
        "nodes": [{"name": names[v], "endpoint": f"http://{names[v]}/api", "inputs": {}} for v in range(n)],
        "edges": [{"from": names[int(u)], "to": names[int(v)]} for u, v in enc.edges[e0:e1]],
    }

"""Row-sharded registry across the GPUs of one box: one process per GPU, ``torch.distributed``.

Partition: rank g owns the contiguous rows
``[g*ceil(N/S), min((g+1)*ceil(N/S), N))``; service ids stay global row numbers (``id_base``).
Queries are replicated; every rank answers them on its shard with the exact local top-k
(``ServiceIndex``), the sorted (score_f64, id) lists are exchanged with ONE all-gather of packed blocks
(S * Q * k * 12 bytes: 6 MiB for config 3 -- latency-bound on NVLink 5 / NVSwitch) and every rank merges
them with the K5 kernel using the same (score, id) key, so the result is identical to the
unsharded search.  There is no other collective on the data path.

The merge is also written for plain CPU tensors + gloo (``merge_gathered_cpu``), which is what the
world_size-2 tests exercise here without a GPU; that helper is host logic (a k-way merge of
sorted lists), not a CPU version of the scan.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np


def shard_bounds(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    per = -(-n_rows // world)
    lo = min(rank * per, n_rows)
    return lo, min(lo + per, n_rows)


def merge_gathered_cpu(scores64: np.ndarray, ids: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """[S,Q,k] sorted per-shard lists -> global (ids, scores64) by (score, id); NaN last; -1 ids are padding."""
    s, q, k = ids.shape
    sc = np.transpose(scores64, (1, 0, 2)).reshape(q, s * k)
    idn = np.transpose(ids, (1, 0, 2)).reshape(q, s * k)
    pad = idn < 0
    nan = np.isnan(sc)
    key_s = np.where(nan | pad, np.inf, sc)
    cls = pad.astype(np.int8) * 2 + (nan & ~pad).astype(np.int8)       # numbers < NaN < padding
    order = np.lexsort((idn, key_s, cls), axis=1)[:, :k]
    out_ids = np.take_along_axis(idn, order, axis=1)
    out_sc = np.take_along_axis(sc, order, axis=1)
    out_sc = np.where(out_ids < 0, np.inf, out_sc)
    return out_ids.astype(np.int32), out_sc


def block_bytes(n_queries: int, k: int) -> int:
    """Bytes of one rank's block in the packed exchange: [Q*k fp64 scores][Q*k int32 ids], padded to 16 (the layout
    ``mcp_merge_topk_packed_device`` reads; the same arithmetic as ``mcp_merge_packed_block_bytes``)."""
    return (n_queries * k * 12 + 15) & ~15


def pack_block_cpu(scores64: np.ndarray, ids: np.ndarray) -> np.ndarray:
    q, k = ids.shape
    blk = np.zeros(block_bytes(q, k), np.uint8)
    blk[: q * k * 8] = np.ascontiguousarray(scores64, np.float64).view(np.uint8).reshape(-1)
    blk[q * k * 8: q * k * 12] = np.ascontiguousarray(ids, np.int32).view(np.uint8).reshape(-1)
    return blk


def unpack_blocks_cpu(packed: np.ndarray, world: int, n_queries: int, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """gathered [world * block] uint8 -> ([S,Q,k] fp64 scores, [S,Q,k] int32 ids)."""
    b = block_bytes(n_queries, k)
    blocks = np.ascontiguousarray(packed, np.uint8).reshape(world, b)
    qk = n_queries * k
    sc = np.stack([blocks[s, : qk * 8].copy().view(np.float64).reshape(n_queries, k) for s in range(world)])
    ids = np.stack([blocks[s, qk * 8: qk * 12].copy().view(np.int32).reshape(n_queries, k) for s in range(world)])
    return sc, ids


class ShardedIndex:
    """Per-rank handle: local ``ServiceIndex`` + the exchange step."""

    def __init__(self, dim: int, metric="cosine", device: Optional[int] = None, group=None):
        import torch.distributed as dist

        from .index import ServiceIndex

        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        import torch

        self.device = torch.cuda.current_device() if device is None else device
        self.local = ServiceIndex(dim, metric, self.device)
        self._gather = None
        self._last = None
        self._host = None

    def adopt_shard(self, rows, n_total: int):
        lo, hi = shard_bounds(n_total, self.world, self.rank)
        assert rows.shape[0] == hi - lo, "shard has the wrong number of rows"
        self.local.adopt(rows, id_base=lo)
        return self

    def load_file_shard(self, path, n_total: Optional[int] = None):
        """Every rank reads only its own row range of the table file (``ServiceIndex.load_file`` window)."""
        from .index import ServiceIndex

        if n_total is None:
            n_total = ServiceIndex.open_table(path)[0].shape[0]
        lo, hi = shard_bounds(n_total, self.world, self.rank)
        self.local.load_file(path, lo, hi - lo)
        return self

    def set_penalty(self, pen):
        self.local.set_penalty(pen)
        return self

    def search_device(self, queries, k: int):
        """-> (ids, scores32, scores64) on this rank's GPU, identical on every rank.  Enqueues only: the local search writes
        its (score fp64, id) lists straight into this rank's block of one packed buffer, ONE all-gather exchanges the blocks
        (12 B x Q x k per rank) and K5 merges them.  ``check()`` tells whether any rank left a query uncertified."""
        import torch

        from .index import merge_topk_packed_device, packed_block_bytes, packed_views

        q = queries.shape[0]
        if self.world == 1:
            self._last = (queries, k, None)
            return self.local.search_device(queries, k)
        blk = packed_block_bytes(q, k)
        if self._gather is None or self._gather[0] != (q, k):
            dev = queries.device
            self._gather = ((q, k), torch.zeros(blk, dtype=torch.uint8, device=dev),
                            torch.empty(self.world * blk, dtype=torch.uint8, device=dev),
                            torch.empty((q, k), dtype=torch.float32, device=dev))
        _, mine, allb, sc32 = self._gather
        ids, sc64 = packed_views(mine, q, k)
        self.local.search_device(queries, k, out=(ids, sc32, sc64))
        self.dist.all_gather_into_tensor(allb, mine, group=self.group)
        self._last = (queries, k, (ids, sc32, sc64))
        return merge_topk_packed_device(allb, self.world, q, k)

    def check(self, repair: bool = True):
        """Blocks until this rank's searches have finished.  -> (certificate failures, unresolved) summed over ranks since the
        previous check.  Unresolved queries (a fallback list overflowed -- rare) hold uncertified rows: with ``repair`` the
        ranks that have some re-answer exactly those with the dense fp64 scan, and the LAST search is exchanged and merged
        again; the repaired result is returned as a third element."""
        import torch

        fails, unres = self.local.sync_status()
        if self.world == 1:
            if unres and repair and self._last is not None and self._last[2] is None:
                raise RuntimeError("unresolved queries on an unsharded search: call ServiceIndex.repair_device with the outputs")
            return fails, unres, None
        t = torch.tensor([fails, unres], dtype=torch.int64, device=torch.device("cuda", self.device))
        self.dist.all_reduce(t, group=self.group)
        tot_f, tot_u = (int(x) for x in t.tolist())
        fixed = None
        if tot_u and repair and self._last is not None:
            from .index import merge_topk_packed_device

            queries, k, out = self._last
            if unres:
                self.local.repair_device(queries, k, out)
            (_q, _k), mine, allb, _ = self._gather
            self.dist.all_gather_into_tensor(allb, mine, group=self.group)
            fixed = merge_topk_packed_device(allb, self.world, queries.shape[0], k)
            torch.cuda.synchronize()
        return tot_f, tot_u, fixed

    def search_host(self, q_host, k: int):
        """Pinned host queries in, pinned host results out, every copy inside: H2D, local search, one all-gather, merge, D2H,
        then ``check`` (a stream synchronisation and a two-word all-reduce of the status).  -> (ids, scores32, scores64) pinned
        CPU tensors, finished and certified, identical on every rank."""
        import torch

        q, dim = q_host.shape
        dev = torch.device("cuda", self.device)
        if self._host is None or self._host[0] != (q, k):
            self._host = ((q, k), torch.empty((q, dim), dtype=torch.float32, device=dev),
                          tuple(torch.empty((q, k), dtype=dt, pin_memory=True) for dt in (torch.int32, torch.float32, torch.float64)))
        _, qd, outs = self._host
        qd.copy_(q_host, non_blocking=True)
        res = self.search_device(qd, k)
        for o, r in zip(outs, res):
            o.copy_(r, non_blocking=True)
        _, unres, fixed = self.check()
        if fixed is not None:
            for o, r in zip(outs, fixed):
                o.copy_(r)
        else:
            torch.cuda.current_stream().synchronize()
        return outs

    def search(self, queries, k: int):
        """search_device + check: finished, certified results."""
        out = self.search_device(queries, k)
        _, unres, fixed = self.check()
        return fixed if fixed is not None else out

    def sync_status(self):
        return self.local.sync_status()

    def close(self):
        self.local.close()

"""Row-sharded registry across the GPUs of one box: one process per GPU, ``torch.distributed``.

Partition: rank g owns the contiguous rows
``[g*ceil(N/S), min((g+1)*ceil(N/S), N))``; service ids stay global row numbers (``id_base``).
Queries are replicated; every rank answers them on its shard with the exact local top-k
(``ServiceIndex``), the sorted (score_f64, id) lists are exchanged with one all-gather
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

    def adopt_shard(self, rows, n_total: int):
        lo, hi = shard_bounds(n_total, self.world, self.rank)
        assert rows.shape[0] == hi - lo, "shard has the wrong number of rows"
        self.local.adopt(rows, id_base=lo)
        return self

    def set_penalty(self, pen):
        self.local.set_penalty(pen)
        return self

    def search_device(self, queries, k: int):
        """-> (ids, scores32, scores64) on this rank's GPU, identical on every rank."""
        import torch

        from .index import merge_topk_device

        ids, sc, sc64 = self.local.search_device(queries, k)
        if self.world == 1:
            return ids, sc, sc64
        q = queries.shape[0]
        if self._gather is None or self._gather[0].shape != (self.world, q, k):
            self._gather = (torch.empty((self.world, q, k), dtype=torch.float64, device=queries.device),
                            torch.empty((self.world, q, k), dtype=torch.int32, device=queries.device))
        g_sc, g_ids = self._gather
        self.dist.all_gather_into_tensor(g_sc.view(-1), sc64.view(-1), group=self.group)
        self.dist.all_gather_into_tensor(g_ids.view(-1), ids.view(-1), group=self.group)
        return merge_topk_device(g_sc, g_ids)

    def sync_status(self):
        return self.local.sync_status()

    def close(self):
        self.local.close()

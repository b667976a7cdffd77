#!/usr/bin/env python
"""Where a sharded step goes (run under torchrun, one rank per GPU): CUDA-event times of the local search, the all-gather
of the packed (score, id) blocks and the K5 merge, per rank, max over ranks.  Config 3 by default.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/sharded_breakdown.py
"""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import sharded  # noqa: E402
from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200.index import (  # noqa: E402
    merge_topk_packed_device, packed_block_bytes, packed_views)

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
sys.stdout.flush()
saved = os.dup(1)
os.dup2(2, 1)
dist.init_process_group("nccl", device_id=dev)
cfg = bench.CONFIGS["config3"]
n, q, k, dim = cfg["n"], cfg["q"], cfg["k"], cfg["dim"]
lo, hi = sharded.shard_bounds(n, world, rank)
rows = bench.make_shard(torch, dev, lo, hi, dim, cfg["seeds"][0] + 7919 * rank)
pen = bench.make_penalty(torch, dev, hi - lo, cfg["seeds"][2] + 7919 * rank)
qs = bench.make_queries(torch, dev, q, dim, cfg["seeds"][1])
sh = sharded.ShardedIndex(dim, "cosine", local)
sh.adopt_shard(rows, n).set_penalty(pen)
for _ in range(30):
    sh.search_device(qs, k)
sh.check()
blk = packed_block_bytes(q, k)
mine = torch.zeros(blk, dtype=torch.uint8, device=dev)
allb = torch.empty(world * blk, dtype=torch.uint8, device=dev)
sc32 = torch.empty((q, k), dtype=torch.float32, device=dev)
ids, sc64 = packed_views(mine, q, k)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
acc = [0.0, 0.0, 0.0]
steps = 50
for _ in range(steps):
    dist.barrier()
    ev[0].record()
    sh.local.search_device(qs, k, out=(ids, sc32, sc64))
    ev[1].record()
    dist.all_gather_into_tensor(allb, mine)
    ev[2].record()
    merge_topk_packed_device(allb, world, q, k)
    ev[3].record()
    torch.cuda.synchronize()
    for i in range(3):
        acc[i] += ev[i].elapsed_time(ev[i + 1])
t = torch.tensor(acc, device=dev) / steps
dist.all_reduce(t, op=dist.ReduceOp.MAX)
sh.check()
os.dup2(saved, 1)
if rank == 0:
    print(json.dumps({"n_gpus": world, "local_search_ms": round(t[0].item(), 4), "all_gather_ms": round(t[1].item(), 4), "merge_ms": round(t[2].item(), 4),
                      "all_gather_bytes_per_rank": blk, "note": "max over ranks of per-rank CUDA-event means; a barrier precedes every step, so the all-gather time excludes waiting for slower ranks' searches only approximately"}))
sh.close()
dist.destroy_process_group()

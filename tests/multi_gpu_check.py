"""Multi-GPU parity check, run under torchrun on a real box (tests/test_gpu_scale.py launches it when the box has two or
more GPUs; by hand:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py)

Every rank builds its row shard of one seeded table, answers the same queries through ShardedIndex
(local exact top-k -> NCCL all-gather -> K5 merge); rank 0 also answers them on the whole table with
one ServiceIndex.  ids must be identical and fp64 scores bit-equal on every rank.
"""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402
from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import sharded  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n, d, q, k = 300_007, 1536, 200, 64
    g = torch.Generator(device=dev).manual_seed(1234)          # same seed everywhere: every rank can build any slice
    full = torch.randn((n, d), generator=g, device=dev)
    full[777] = full[5]                                        # a tie that straddles nothing; and one across shards:
    full[n - 3] = full[11]
    pen = torch.rand(n, generator=g, device=dev) * 0.02
    pen[777], pen[n - 3] = pen[5], pen[11]
    qs = torch.randn((q, d), generator=g, device=dev)
    qs[0] = full[11]
    lo, hi = sharded.shard_bounds(n, world, rank)
    sh = sharded.ShardedIndex(d, "cosine", local)
    sh.adopt_shard(full[lo:hi], n).set_penalty(pen[lo:hi].contiguous())
    ok = True
    for path in ("auto", "gemm_tf32", "stream_fp32"):
        sh.local.set_option("path", path)
        ids, s32, s64 = sh.search_device(qs, k)
        assert sh.check()[1] == 0
        one = pkg.ServiceIndex(d, "cosine", local).adopt(full)
        one.set_penalty(pen)
        wi, w32, w64 = one.search_device(qs, k)
        assert one.sync_status()[1] == 0
        torch.cuda.synchronize()
        same = torch.equal(ids, wi) and torch.equal(s64, w64)
        tie = ids[0, 0].item() == 11 and ids[0, 1].item() == n - 3
        ok = ok and same and tie
        one.close()
    t = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MULTI_GPU_CHECK", "PASS" if t.item() == 1 else "FAIL", f"world={world}")
    sh.close()
    dist.destroy_process_group()
    sys.exit(0 if t.item() == 1 else 1)


if __name__ == "__main__":
    main()

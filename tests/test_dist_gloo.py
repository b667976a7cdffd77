"""The N > 1 exchange step on CPU: world_size 2, gloo.  Each rank holds a row shard, answers the
same queries on it (the oracle stands in for the local GPU search -- this is a test of the
sharding / all-gather / merge logic, not of the scan), gathers and merges."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    import oracle
    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import sharded

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c = cases.config2_small(oracle.telemetry_penalty, n=1500, q=7, k=24)
        lo, hi = sharded.shard_bounds(1500, world, rank)
        ids, s64, _ = oracle.topk(c["rows"][lo:hi], c["queries"], c["k"], 0, c["penalty"][lo:hi], id_base=lo, threads=1)
        g_sc = torch.empty((world, 7, 24), dtype=torch.float64)
        g_id = torch.empty((world, 7, 24), dtype=torch.int32)
        dist.all_gather_into_tensor(g_sc.view(-1), torch.from_numpy(s64).view(-1))
        dist.all_gather_into_tensor(g_id.view(-1), torch.from_numpy(ids).view(-1))
        mi, ms = sharded.merge_gathered_cpu(g_sc.numpy(), g_id.numpy())
        gi, gs, _ = oracle.topk(c["rows"], c["queries"], c["k"], 0, c["penalty"], threads=1)
        ok = np.array_equal(mi, gi) and np.array_equal(ms, gs)
        # the packed exchange the GPU path uses: one all-gather of [scores | ids] blocks
        blk = torch.from_numpy(sharded.pack_block_cpu(s64, ids))
        allb = torch.empty(world * blk.numel(), dtype=torch.uint8)
        dist.all_gather_into_tensor(allb, blk)
        p_sc, p_id = sharded.unpack_blocks_cpu(allb.numpy(), world, 7, 24)
        pi, ps = sharded.merge_gathered_cpu(p_sc, p_id)
        ok = ok and np.array_equal(pi, gi) and np.array_equal(ps, gs) and blk.numel() == sharded.block_bytes(7, 24)
        t = torch.tensor([int(ok)])
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if rank == 0:
            ret.put(int(t.item()))
    finally:
        dist.destroy_process_group()


def test_two_rank_shard_gather_merge():
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) == 1

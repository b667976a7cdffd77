#!/usr/bin/env python
"""K6 timing probe: n synthetic plan DAGs (config 4 generator), with and without the topological_sort order, through each
kernel choice (MCP_DAG_KERNEL = warp | fifo256 | fifo | fifo1024; unset = the library's own choice), and a bit-for-bit
comparison of every output of the thread-per-DAG kernels against the warp kernel's."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dup = len(sys.argv) > 2 and sys.argv[2] == "dup"
dev = torch.device("cuda:0")
enc = pkg.dag.synthetic_dags(n, seed=4004)
edges = enc.edges
edge_off = enc.edge_off
if dup:      # every edge list twice over (repeated edges must collapse), the second copy reversed in order
    parts, off = [], [0]
    for i in range(n):
        e = edges[edge_off[i]:edge_off[i + 1]]
        parts.append(e); parts.append(e[::-1])
        off.append(off[-1] + 2 * len(e))
    edges = np.concatenate(parts) if parts else edges
    edge_off = np.asarray(off, dtype=np.uint32)
nn = torch.from_numpy(enc.n_nodes).to(dev)
off = torch.from_numpy(edge_off.view(np.int32)).to(dev)
ed = torch.from_numpy(np.ascontiguousarray(edges)).to(dev).contiguous()
ref = {}
for choice in ("warp", "fifo256", "fifo", "fifo1024", ""):
    if choice:
        os.environ["MCP_DAG_KERNEL"] = choice
    else:
        os.environ.pop("MCP_DAG_KERNEL", None)
    for order in (True, False):
        out = pkg.dag.levels_device(nn, off, ed, with_order=order)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            pkg.dag.levels_device(nn, off, ed, out=out, with_order=order)
        e1.record()
        torch.cuda.synchronize()
        got = [None if o is None else o.cpu().numpy() for o in out]
        if choice == "warp":
            ref[order] = got
            same = "reference"
        else:
            same = all((a is None and b is None) or np.array_equal(a, b) for a, b in zip(got, ref[order]))
        print(f"{n} DAGs, kernel={choice or 'auto':14s} {'with order ' if order else 'levels only'}: "
              f"{e0.elapsed_time(e1) / 5:.3f} ms   equal to warp kernel: {same}", flush=True)

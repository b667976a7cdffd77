#!/usr/bin/env python
"""K6 timing probe: synthetic plan DAGs (config 4 generator), with and without the topological_sort order, through each
kernel (MCP_DAG_KERNEL = warp | fifo, MCP_DAG_SUB = DAGs a CTA sorts at a time; unset = the library's own choice), at
several batch sizes, and a bit-for-bit comparison of every output of the thread-per-DAG kernel against the warp kernel's.

  python tools/dag_probe.py [n_dags] [dup|tiny]   dup: every edge list repeated (reversed) -- repeated edges must collapse;
                                                  tiny: plans of 2..10 nodes
"""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402

n_max = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dup = len(sys.argv) > 2 and sys.argv[2] == "dup"
tiny = len(sys.argv) > 2 and sys.argv[2] == "tiny"
dev = torch.device("cuda:0")
enc = pkg.dag.synthetic_dags(n_max, seed=4004)
edges = enc.edges
edge_off = enc.edge_off
if tiny:     # plans the size a planner emits: 2..10 nodes, every forward pair an edge with probability 0.3
    rng = np.random.default_rng(77)
    nn_t = rng.integers(2, 11, n_max).astype(np.uint8)
    iu, ju = np.triu_indices(10, 1)
    keep = (rng.random((n_max, iu.size)) < 0.3) & (ju[None, :] < nn_t[:, None])
    cnt = keep.sum(1)
    rows, cols = np.nonzero(keep)
    edges = np.stack([iu[cols], ju[cols]], axis=1).astype(np.uint8)
    edge_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint32)
    enc = pkg.dag.EncodedDags(nn_t, edge_off, edges, [])
if dup:
    parts, off = [], [0]
    for i in range(n_max):
        e = edges[edge_off[i]:edge_off[i + 1]]
        parts.append(e); parts.append(e[::-1])
        off.append(off[-1] + 2 * len(e))
    edges = np.concatenate(parts) if parts else edges
    edge_off = np.asarray(off, dtype=np.uint32)
nn_all = torch.from_numpy(enc.n_nodes).to(dev)
off_all = torch.from_numpy(edge_off.view(np.int32)).to(dev)
ed = torch.from_numpy(np.ascontiguousarray(edges)).to(dev).contiguous()
sizes = [n for n in (1024, 4096, 16384, 50_000, 100_000, 200_000, 400_000) if n < n_max] + [n_max]
choices = [("warp", None), ("fifo", "1024"), ("fifo", None), (None, None)]
if os.environ.get("DAG_PROBE_ONLY"):        # one size, the library's own choice after the reference: for an ncu capture
    sizes, choices = [n_max], [("warp", None), (None, None)]
for n in sizes:
    nn, off = nn_all[:n], off_all[:n + 1]
    ref = {}
    for kernel, sub in choices:
        for key, val in (("MCP_DAG_KERNEL", kernel), ("MCP_DAG_SUB", sub)):
            if val is None:
                os.environ.pop(key, None)
            else:
                os.environ[key] = val
        line = f"{n:8d} DAGs  kernel={kernel or 'auto':5s} sub={sub or '-':5s}"
        for order in (True, False):
            out = pkg.dag.levels_device(nn, off, ed, with_order=order)
            torch.cuda.synchronize()
            reps = 5 if n >= 200_000 else 20
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                pkg.dag.levels_device(nn, off, ed, out=out, with_order=order)
            e1.record()
            torch.cuda.synchronize()
            got = [None if o is None else o.cpu().numpy() for o in out]
            if kernel == "warp":
                ref[order] = got
                same = "ref"
            else:
                same = all((a is None and b is None) or np.array_equal(a, b) for a, b in zip(got, ref[order]))
            line += f"   {'order ' if order else 'levels'} {e0.elapsed_time(e1) / reps:7.3f} ms (equal: {same})"
        print(line, flush=True)

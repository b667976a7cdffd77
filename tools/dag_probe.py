#!/usr/bin/env python
"""K6 timing probe: n synthetic plan DAGs (config 4 generator), with and without the topological_sort order."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dev = torch.device("cuda:0")
enc = pkg.dag.synthetic_dags(n, seed=4004)
nn = torch.from_numpy(enc.n_nodes).to(dev)
off = torch.from_numpy(enc.edge_off.view(np.int32)).to(dev)
ed = torch.from_numpy(enc.edges).to(dev).contiguous()
for order in (True, False):
    out = pkg.dag.levels_device(nn, off, ed, with_order=order)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        pkg.dag.levels_device(nn, off, ed, out=out, with_order=order)
    e1.record()
    torch.cuda.synchronize()
    print(f"{n} DAGs, {'with order' if order else 'levels only'}: {e0.elapsed_time(e1) / 5:.3f} ms")

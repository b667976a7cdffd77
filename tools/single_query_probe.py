#!/usr/bin/env python
"""Where a single-intent search goes (config 5's shape: 1 M services x 1536, top-8): device-resident search by CUDA events,
the host-buffer call by wall clock, and the coarse scan alone (option time_coarse)."""
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as mcp  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
d, k = 1536, 8
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
rows = torch.randn((n, d), device=dev, generator=g)
idx = mcp.ServiceIndex(d, "cosine")
idx.adopt(rows)
idx.set_penalty(torch.rand(n, device=dev, generator=g) * 0.05)
q = torch.randn((1, d), device=dev, generator=g)
out = idx.search_device(q, k)
for _ in range(10):
    idx.search_device(q, k, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    idx.search_device(q, k, out=out)
e1.record()
torch.cuda.synchronize()
print(f"search_device, back to back: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per search")
idx.set_option("time_coarse", 1)
s0 = idx.stats()
for _ in range(50):
    idx.search_device(q, k, out=out)
torch.cuda.synchronize()
s1 = idx.stats()
idx.set_option("time_coarse", 0)
print(f"coarse scan alone: {(s1['coarse_ms'] - s0['coarse_ms']) / max(1, s1['coarse_timed'] - s0['coarse_timed']) * 1e3:.1f} us; "
      f"launches per search {(s1['kernel_launches'] - s0['kernel_launches']) / 50:.1f}; path {s1['last_path']} k' {s1['last_kprime']}")
qh = q.cpu().numpy()
for _ in range(10):
    idx.search(qh, k)
ts = []
for _ in range(reps):
    t0 = time.perf_counter()
    idx.search(qh, k)
    ts.append(time.perf_counter() - t0)
ts = np.sort(np.asarray(ts)) * 1e6
print(f"mcp_index_search (host buffers): p50 {ts[len(ts) // 2]:.1f} us, p99 {ts[int(len(ts) * 0.99)]:.1f} us; "
      f"shadow bytes / 6551 GB/s = {n * (d * 2 + 12) / 6551e9 * 1e6:.1f} us")

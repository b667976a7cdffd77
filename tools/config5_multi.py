#!/usr/bin/env python
"""Config 5 through ONE handle over all GPUs of the box (mcp_index_create_multi): single-intent latency over 1M services,
host buffers in and out (the call GraphPlanner.retrieve makes), beside the same on one GPU.

  python tools/config5_multi.py [--rows 1000000] [--requests 2000]
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000)
ap.add_argument("--requests", type=int, default=2000)
ap.add_argument("--k", type=int, default=8)
a = ap.parse_args()
dim = 1536
g = torch.Generator(device="cuda:0").manual_seed(1005)
rows_d = torch.randn((a.rows, dim), generator=g, device="cuda:0")
rows_d /= rows_d.norm(dim=1, keepdim=True)
rows = rows_d.cpu().numpy()
del rows_d
qs = np.random.default_rng(7).standard_normal((a.requests, dim), dtype=np.float32)
out = {"rows": a.rows, "k": a.k}
ngpu = torch.cuda.device_count()
for devs in ([0], list(range(ngpu))) if ngpu > 1 else ([0], [0, 0]):
    with pkg.ServiceIndex(dim, "cosine", devs if len(devs) > 1 else devs[0]) as idx:
        t0 = time.perf_counter()
        idx.load(rows)
        t_load = time.perf_counter() - t0
        for i in range(50):
            idx.search(qs[i:i + 1], a.k)
        lat = []
        for i in range(a.requests):
            t0 = time.perf_counter()
            ids, _, _ = idx.search(qs[i:i + 1], a.k)
            lat.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        idx.search(qs[:1024], 64)
        t_b = time.perf_counter() - t0
        t0 = time.perf_counter()
        idx.search(qs[:1024], 64)
        t_b = min(t_b, time.perf_counter() - t0)
        # the same through the reference-facing call: GraphPlanner.retrieve (hash embedding + search + name lookup)
        from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

        class _Reg:
            def list_services(self):
                return []

        planner = cp.GraphPlanner(_Reg(), top_k=a.k)
        planner.attach_index(idx, [f"svc{i:07d}" for i in range(a.rows)])
        intents = [f"intent number {i}: book, translate, summarise and notify" for i in range(a.requests)]
        for t in intents[:50]:
            planner.retrieve(t)
        lat_p = []
        for t in intents:
            t0 = time.perf_counter()
            planner.retrieve(t)
            lat_p.append(time.perf_counter() - t0)
        planner._index = None
        out[f"devices_{len(devs)}" + ("_same_gpu" if len(set(devs)) < len(devs) else "")] = {
            "planner_retrieve_p50_ms": round(float(np.percentile(lat_p, 50)) * 1e3, 4), "planner_retrieve_p99_ms": round(float(np.percentile(lat_p, 99)) * 1e3, 4),
            "p50_ms": round(float(np.percentile(lat, 50)) * 1e3, 4), "p99_ms": round(float(np.percentile(lat, 99)) * 1e3, 4),
            "batch_1024_top64_ms": round(t_b * 1e3, 3), "load_s": round(t_load, 2), "first_ids": ids[0, :3].tolist()}
print(json.dumps(out))

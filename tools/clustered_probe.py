#!/usr/bin/env python
"""Clustered-registry probe: search time and certificate failures at several oversampling settings, and the fp64 scan's own
throughput (path = exact_fp64, dense mode) -- what the fallback costs when it runs.

  python tools/clustered_probe.py
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402
import bench  # noqa: E402

dev = torch.device("cuda:0")
n, dim, k = 100_000, 1536, 32
rc, qc = bench.make_clustered(torch, dev, n, dim, 77)
rr = bench.make_shard(torch, dev, 0, n, dim, 5)
qr = bench.make_queries(torch, dev, 256, dim, 6)


def run(name, rows, qs, **opts):
    idx = pkg.ServiceIndex(dim, "cosine", 0).adopt(rows)
    for o, v in opts.items():
        idx.set_option(o, v)
    for _ in range(30):
        idx.search_device(qs, k)
    idx.sync_status()
    ms = bench.time_steps(torch, lambda: idx.search_device(qs, k), 100)
    f, u = idx.sync_status()
    print(f"{name:28s} {opts}: {ms * 1e3:7.1f} us per batch of {qs.shape[0]}, failures/batch {f / 100:.2f}, unresolved {u}, k' {idx.stats()['last_kprime']}", flush=True)
    idx.close()


run("random", rr, qr)
run("clustered", rc, qc)
for kp in (64, 128, 256):
    run("clustered", rc, qc, kprime=kp)
run("clustered tf32", rc, qc, path=2)
for nq in (1, 4, 8, 16):
    run("fp64 scan (dense), random", rr, qr[:nq].contiguous(), path=3)

#!/usr/bin/env python
"""BASELINE config 5: end-to-end /plan_and_execute with a stubbed LLM and mocked HTTP services over a
1M-service registry; retrieval p50 / p99 latency.

  python tools/config5_e2e.py [--rows 1000000] [--requests 10000] [--clients 32] [--k 8]

Everything around the hot path is a stand-in (dict-backed Redis, canned LLM answer, httpx.MockTransport);
the retrieval itself is the real thing: GraphPlanner.retrieve -> ServiceIndex.search -> mcp_index_search on the GPU
with host buffers (one intent per request, K1 fp32 streaming path).  Prints one JSON object.
"""
import argparse
import json
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


class LazyRedis:
    """1M service records without 1M Python dicts: records are synthesised on GET."""

    def __init__(self, n, cost):
        self.n, self.cost = n, cost

    def scan_iter(self, pattern):
        return (f"mcp:service:svc{i:07d}" for i in range(self.n))

    def get(self, key):
        if key.startswith("mcp:service:svc"):
            i = int(key[len("mcp:service:svc"):])
            return json.dumps({"name": f"svc{i:07d}", "endpoint": f"http://svc{i:07d}/api", "input_schema": {"x": "str"},
                               "output_schema": {"y": "str"}, "cost_profile": float(self.cost[i]), "fallback": None})
        return None


def pct(a, p):
    return float(np.percentile(np.asarray(a) * 1e3, p))


def run(pkg, rows, pen, requests=10_000, clients=32, k=8, endpoint_requests=300):
    """rows / pen: CUDA tensors (the table is adopted in place).  -> dict of latencies (milliseconds)."""
    import httpx
    from fastapi.testclient import TestClient

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    n = rows.shape[0]
    idx = pkg.ServiceIndex(rows.shape[1], "cosine", rows.device.index or 0).adopt(rows)
    idx.set_penalty(pen)
    names = [f"svc{i:07d}" for i in range(n)]
    redis = LazyRedis(n, np.full(n, 0.005, np.float32))

    def llm(prompt):
        svc = [ln.split(" ")[1] for ln in prompt.splitlines() if ln.startswith("- ")][:3]
        nodes = [{"name": s, "endpoint": f"http://{s}/api", "inputs": {}} for s in svc]
        edges = [{"from": a, "to": b} for a, b in zip(svc, svc[1:])]
        return json.dumps({"nodes": nodes, "edges": edges})

    planner = cp.GraphPlanner(cp.ServiceRegistry(client=redis), llm=llm, top_k=k)
    planner.attach_index(idx, names)
    orch = cp.Orchestrator(client=httpx.AsyncClient(transport=httpx.MockTransport(lambda r: httpx.Response(200, json={"ok": True}))))
    app = cp.create_app(registry=planner.registry, planner=planner, orch=orch)

    intents = [f"intent number {i}: book, translate, summarise and notify" for i in range(requests)]
    for t in intents[:50]:
        planner.retrieve(t)
    # (a) sequential retrieval calls
    lat = []
    for t in intents:
        t0 = time.perf_counter()
        planner.retrieve(t)
        lat.append(time.perf_counter() - t0)
    # search only (pre-embedded queries): what the C ABI call itself costs
    q = np.stack([cp.hash_embedding(t) for t in intents[:2000]])
    lat_abi = []
    for i in range(q.shape[0]):
        t0 = time.perf_counter()
        idx.search(q[i:i + 1], k)
        lat_abi.append(time.perf_counter() - t0)
    # (b) concurrent clients (threads; the index serialises calls on its lock, the GIL is released inside)
    lat_c, lock = [], threading.Lock()

    def worker(chunk):
        loc = []
        for t in chunk:
            t0 = time.perf_counter()
            planner.retrieve(t)
            loc.append(time.perf_counter() - t0)
        with lock:
            lat_c.extend(loc)

    per = max(1, len(intents) // clients)
    threads = [threading.Thread(target=worker, args=(intents[i * per:(i + 1) * per],)) for i in range(clients)]
    t0 = time.perf_counter()
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    wall_c = time.perf_counter() - t0
    # (b2) the same clients through the micro-batcher: concurrent requests share launches
    planner_mb = cp.GraphPlanner(cp.ServiceRegistry(client=redis), llm=llm, top_k=k, micro_batch=True)
    planner_mb.attach_index(idx, names)
    lat_m = []

    def worker_mb(chunk):
        loc = []
        for t in chunk:
            t0 = time.perf_counter()
            planner_mb.retrieve(t)
            loc.append(time.perf_counter() - t0)
        with lock:
            lat_m.extend(loc)

    threads = [threading.Thread(target=worker_mb, args=(intents[i * per:(i + 1) * per],)) for i in range(clients)]
    t0 = time.perf_counter()
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    wall_m = time.perf_counter() - t0
    # (c) batched retrieval of the same intents (what request micro-batching would reach)
    t0 = time.perf_counter()
    planner.retrieve_batch(intents[:1024], k)
    batch_1024 = time.perf_counter() - t0
    # (d) the endpoint, end to end
    client = TestClient(app, raise_server_exceptions=True)
    lat_e = []
    for t in intents[:endpoint_requests]:
        t0 = time.perf_counter()
        r = client.post("/plan_and_execute", json={"intent": t})
        lat_e.append(time.perf_counter() - t0)
        assert r.status_code == 200 and len(r.json()["results"]) == 3
    out = {
        "config": f"{n} services x {rows.shape[1]}-d, top-{k} cosine + telemetry, single-intent requests",
        "retrieve_sequential_ms": {"p50": pct(lat, 50), "p99": pct(lat, 99), "mean": float(np.mean(lat) * 1e3), "n": len(lat),
                                   "note": "GraphPlanner.retrieve: hash embedding (numpy) + mcp_index_search + name lookup"},
        "mcp_index_search_ms": {"p50": pct(lat_abi, 50), "p99": pct(lat_abi, 99), "n": len(lat_abi),
                                "hbm_floor_ms_fp32_rows": n * rows.shape[1] * 4 / 6551.4e9 * 1e3,
                                "hbm_floor_ms_fp16_shadow": n * rows.shape[1] * 2 / 6551.4e9 * 1e3, "path": idx.stats()["last_path"]},
        "retrieve_concurrent_ms": {"clients": clients, "p50": pct(lat_c, 50), "p99": pct(lat_c, 99),
                                   "throughput_rps": len(lat_c) / wall_c},
        "retrieve_concurrent_microbatched_ms": {"clients": clients, "p50": pct(lat_m, 50), "p99": pct(lat_m, 99),
                                                "throughput_rps": len(lat_m) / wall_m, "launches": planner_mb._batcher.launches,
                                                "requests": planner_mb._batcher.requests},
        "retrieve_batch_1024_ms": batch_1024 * 1e3,
        "plan_and_execute_ms": {"p50": pct(lat_e, 50), "p99": pct(lat_e, 99), "n": len(lat_e),
                                "note": "FastAPI TestClient, stub LLM, 3 mocked services per plan"},
    }
    if planner_mb._batcher is not None:
        planner_mb._batcher.close()
    idx.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--requests", type=int, default=10_000)
    ap.add_argument("--clients", type=int, default=32)
    ap.add_argument("--k", type=int, default=8)
    args = ap.parse_args()

    import torch

    import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg
    import bench

    dev = torch.device("cuda:0")
    rows = bench.make_shard(torch, dev, 0, args.rows, 1536, 1005)
    pen = bench.make_penalty(torch, dev, args.rows, 2005)
    print(json.dumps(run(pkg, rows, pen, args.requests, args.clients, args.k)))


if __name__ == "__main__":
    main()

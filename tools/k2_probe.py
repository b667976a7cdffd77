#!/usr/bin/env python
"""K2 timing probe: one table, many option sets.  Prints the coarse kernel's CUDA-event time (option time_coarse)
and the whole-search time per set.  Sets whose gemm_debug switches parts of the kernel off give WRONG results;
they are timing experiments (what bounds K2: loads, MMAs, epilogue).

  python tools/k2_probe.py --rows 1250000 --queries 1024 --k 64 --path gemm_f16 \
         --sets "" "gemm_sync=0" "gemm_sync=2" "gemm_debug=1" "gemm_debug=6" "gemm_debug=8"
"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_250_000)
ap.add_argument("--queries", type=int, default=1024)
ap.add_argument("--k", type=int, default=64)
ap.add_argument("--dim", type=int, default=1536)
ap.add_argument("--path", default="gemm_f16")
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--sets", nargs="*", default=[""])
a = ap.parse_args()
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1003)
rows = torch.empty((a.rows, a.dim), device=dev)
for lo in range(0, a.rows, 262144):
    hi = min(a.rows, lo + 262144)
    torch.randn((hi - lo, a.dim), generator=g, device=dev, out=rows[lo:hi])
    rows[lo:hi] /= rows[lo:hi].norm(dim=1, keepdim=True)
qs = torch.randn((a.queries, a.dim), generator=g, device=dev)
qs /= qs.norm(dim=1, keepdim=True)
for spec in a.sets:
    idx = pkg.ServiceIndex(a.dim, "cosine", 0).adopt(rows)
    idx.set_option("path", a.path).set_option("time_coarse", 1)
    for kv in spec.split(","):
        if kv:
            name, v = kv.split("=")
            idx.set_option(name, int(v))
    for _ in range(3):
        idx.search_device(qs, a.k)
    idx.sync_status()
    s0 = idx.stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        idx.search_device(qs, a.k)
    e1.record()
    torch.cuda.synchronize()
    fails, unres = idx.sync_status()
    s1 = idx.stats()
    k2 = (s1["coarse_ms"] - s0["coarse_ms"]) / max(1, s1["coarse_timed"] - s0["coarse_timed"])
    print(f"{a.path} [{spec or 'default'}]: K2 {k2:.4f} ms, search {e0.elapsed_time(e1) / a.steps:.4f} ms, "
          f"cert_fail {fails} unresolved {unres} kprime {s1['last_kprime']}", flush=True)
    idx.close()

"""Randomised differential test of the CUDA retrieval path against the CPU oracle (run on a B200):

    python tools/fuzz_parity.py [n_cases] [seed]

Shapes are drawn so that every K2 shape (single CTA, multicast cluster, CTA pair), the bootstrap with and without the early
accumulator hand-back, the K1 path and query chunking all come up; every case must be bit-exact (ids, fp64 scores).
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as mcp  # noqa: E402
import oracle  # noqa: E402


def main():
    n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 12345)
    bad = 0
    t0 = time.time()
    for case in range(n_cases):
        d = int(rng.choice([4, 32, 64, 100, 128, 200, 256]))
        n = int(rng.choice([1, 37, 300, 5000, 66_000, 70_001, 120_000]))
        q = int(rng.choice([1, 7, 8, 33, 128, 129, 256, 300, 512]))
        k = int(rng.choice([1, 8, 32, 64, 200]))
        metric = int(rng.integers(0, 2))
        path = str(rng.choice(["auto", "auto", "gemm_tf32", "gemm_f16", "stream_fp32"]))
        rows = rng.standard_normal((n, d), dtype=np.float32)
        if rng.random() < 0.3 and n > 50:                       # clusters of near-duplicates and exact duplicates
            rows[n // 2:n // 2 + 20] = rows[0] + rng.standard_normal((20, d)).astype(np.float32) * 1e-3
            rows[5] = rows[3]
        qs = rng.standard_normal((q, d), dtype=np.float32)
        if rng.random() < 0.3 and n > 40:
            qs[0] = rows[min(n - 1, 17)]                         # a query that is a row: many rows of the first tile may matter
            rows[:min(n, 40)] += (qs[0] * 0.5).astype(np.float32)
        pen = (rng.random(n) * 0.05).astype(np.float32) if rng.random() < 0.6 else None
        opts = {}
        if rng.random() < 0.3:
            opts["gemm_pair"] = 0
        if rng.random() < 0.2:
            opts["gemm_balance"] = 1
        if rng.random() < 0.15:
            opts["gemm_cluster"] = 1
        want_ids, want_s64, _ = oracle.topk(rows, qs, k, metric, pen)
        with mcp.ServiceIndex(d, "cosine" if metric == 0 else "l2") as idx:
            idx.load(rows)
            if pen is not None:
                idx.set_penalty(pen)
            idx.set_option("path", path)
            for name, v in opts.items():
                idx.set_option(name, v)
            ids, s32, s64 = idx.search(qs, k)
            st = idx.stats()
        ok = np.array_equal(ids, want_ids) and np.array_equal(s64, want_s64, equal_nan=True)
        bad += 0 if ok else 1
        print(f"case {case:3d} n={n:6d} d={d:3d} q={q:3d} k={k:3d} metric={metric} path={path:11s} opts={opts} "
              f"last_path={st['last_path']} cert_fail={st['certificate_failures']} -> {'ok' if ok else 'MISMATCH'}", flush=True)
    print(f"{n_cases - bad}/{n_cases} cases bit-exact in {time.time() - t0:.1f} s")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()

"""Parity at the sizes BASELINE.json quotes, and on data that is not isotropic noise.

  * config 3's per-GPU share at 8 GPUs (1.25 M rows x 1536, 1024 queries, top-64, default options, so every code path that
    only switches on at this size is live: 4 CTA pairs per group, the tile drift limit, early bootstrap hand-back, the
    shared-threshold refresh): every query against the GPU fp64 scan, a sample against the CPU oracle, cosine and L2;
  * a clustered registry (services with near-identical schemas: tight clusters, exact and near-exact duplicates), where the
    gap between rank k and rank k' collapses: still bit-exact, nothing unresolved, and the failure rate is reported;
  * the row-sharded search on 2+ GPUs equals the unsharded one (skips below two GPUs).
"""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def clustered_rows(n, d, seed, n_centres=2000, dup_frac=0.01):
    """Cluster centres on the unit sphere, members = centre + sigma * noise / sqrt(d) with sigma ~ U(0.02, 0.1) per cluster
    (so members sit 0.02 .. 0.1 away from their centre), then 1 % of the rows overwritten with exact copies of other rows and
    with copies perturbed in the last bits.  Returns float32 [n, d], un-normalised."""
    rng = np.random.default_rng(seed)
    centres = rng.standard_normal((n_centres, d)).astype(np.float32)
    centres /= np.linalg.norm(centres, axis=1, keepdims=True)
    which = rng.integers(0, n_centres, n)
    sigma = rng.uniform(0.02, 0.1, n_centres).astype(np.float32)[which, None]
    rows = centres[which] + sigma * rng.standard_normal((n, d)).astype(np.float32) / np.float32(np.sqrt(d))
    n_dup = max(2, int(n * dup_frac))
    src = rng.integers(0, n, n_dup)
    dst = rng.integers(0, n, n_dup)
    rows[dst[: n_dup // 2]] = rows[src[: n_dup // 2]]                                   # exact duplicates: ties -> lowest id
    rows[dst[n_dup // 2:]] = rows[src[n_dup // 2:]] * (1.0 + rng.standard_normal((n_dup - n_dup // 2, d)).astype(np.float32) * 1e-6)
    return rows.astype(np.float32), centres


@pytest.mark.parametrize("metric", ["cosine", "l2"])
def test_config3_share_full_size(mcp, oracle, metric):
    import torch

    dev = torch.device("cuda:0")
    n, d, q, k = 1_250_000, 1536, 1024, 64
    g = torch.Generator(device=dev).manual_seed(1003)
    rows = torch.empty((n, d), device=dev)
    for lo in range(0, n, 250_000):
        torch.randn((250_000, d), generator=g, device=dev, out=rows[lo:lo + 250_000])
        rows[lo:lo + 250_000] /= rows[lo:lo + 250_000].norm(dim=1, keepdim=True)
    qs = torch.randn((q, d), generator=g, device=dev)
    qs /= qs.norm(dim=1, keepdim=True)
    pen = torch.rand(n, generator=g, device=dev) * 0.02
    idx = mcp.ServiceIndex(d, metric).adopt(rows)
    idx.set_penalty(pen)
    ids, s32, s64 = (t.clone() for t in idx.search_device(qs, k))            # default options: path auto
    fails, unresolved = idx.sync_status()
    st = idx.stats()
    assert unresolved == 0
    assert st["last_path"] == 4 and st["last_kprime"] == 96, st              # the fp16 shadow scan, k' = 1.5 k
    assert bool((s64[:, 1:] >= s64[:, :-1]).all())
    for path in ("gemm_tf32", "exact_fp64"):
        idx.set_option("path", path)
        ids_x, _, s64_x = idx.search_device(qs, k)
        assert idx.sync_status()[1] == 0
        assert torch.equal(ids, ids_x) and torch.equal(s64, s64_x), path     # every query == the fp64 scan of every row
    sample = list(range(0, q, 64)) + [q - 1]                                 # 17 queries, every query block
    rows_np = rows.cpu().numpy()
    want_ids, want_s64, _ = oracle.topk(rows_np, qs.cpu().numpy()[sample], k, 0 if metric == "cosine" else 1, pen.cpu().numpy())
    assert np.array_equal(ids.cpu().numpy()[sample], want_ids)
    assert np.array_equal(s64.cpu().numpy()[sample], want_s64)
    assert np.array_equal(s32.cpu().numpy()[sample], want_s64.astype(np.float32))
    assert fails <= 8, f"{fails} of {q} queries failed the certificate on random data: the bound or k' is too tight"
    idx.close()


@pytest.mark.parametrize("path", ["auto", "gemm_tf32", "stream_fp32"])
def test_clustered_registry_stays_exact(mcp, oracle, path):
    import torch

    dev = torch.device("cuda:0")
    n, d, q, k = 100_000, 512, 256, 32
    rows_np, centres = clustered_rows(n, d, 77)
    rng = np.random.default_rng(78)
    # queries: near cluster centres (their whole top-k sits in one tight cluster), copies of rows, and a few random ones
    qs_np = np.concatenate([centres[rng.integers(0, len(centres), q - 64)] + 0.01 * rng.standard_normal((q - 64, d)).astype(np.float32),
                            rows_np[rng.integers(0, n, 32)],
                            rng.standard_normal((32, d)).astype(np.float32)]).astype(np.float32)
    if path == "stream_fp32":
        qs_np = qs_np[:24]
    idx = mcp.ServiceIndex(d, "cosine").adopt(torch.from_numpy(rows_np).to(dev))
    idx.set_option("path", path)
    qs = torch.from_numpy(qs_np).to(dev)
    ids, s32, s64 = (t.clone() for t in idx.search_device(qs, k))
    fails, unresolved = idx.sync_status()
    assert unresolved == 0
    idx.set_option("path", "exact_fp64")
    ids_x, _, s64_x = idx.search_device(qs, k)
    assert idx.sync_status()[1] == 0
    assert torch.equal(ids, ids_x) and torch.equal(s64, s64_x)
    sample = list(range(0, len(qs_np), max(1, len(qs_np) // 16)))
    want_ids, want_s64, _ = oracle.topk(rows_np, qs_np[sample], k, 0, None)
    assert np.array_equal(ids.cpu().numpy()[sample], want_ids) and np.array_equal(s64.cpu().numpy()[sample], want_s64)
    print(f"clustered registry, path {path}: {fails} of {len(qs_np)} queries needed the fp64 scan")
    # host-buffer entry point on the same data: finished, certified answers in one call
    hi, _, hs64 = idx.set_option("path", path).search(qs_np, k)
    assert np.array_equal(hi, ids.cpu().numpy()) and np.array_equal(hs64, s64.cpu().numpy())
    idx.close()


def test_fallback_list_overflow_is_repaired(mcp, oracle):
    """More rows inside the coarse error band of the k-th score than a fallback list holds (3000 near-copies of the query's
    best match, k' = 32): the device entry point must report the query unresolved and repair_device must fix exactly it;
    the host entry point must do both by itself."""
    import torch

    dev = torch.device("cuda:0")
    rng = np.random.default_rng(5)
    d, k = 256, 20
    base = rng.standard_normal(d).astype(np.float32)
    rows = rng.standard_normal((20_000, d)).astype(np.float32)
    rows[1000:4000] = base * (1.0 + (rng.standard_normal((3000, d)) * 1e-7).astype(np.float32))
    qs = np.stack([base, rng.standard_normal(d).astype(np.float32)]).astype(np.float32)
    qs = np.repeat(qs, 8, axis=0)                                           # 16 queries: the tensor path
    want_ids, want_s64, _ = oracle.topk(rows, qs, k, 0, None)
    with mcp.ServiceIndex(d, "cosine") as idx:
        idx.load(rows)
        idx.set_option("kprime", 32)
        got = idx.search(qs, k)
        assert np.array_equal(got[0], want_ids) and np.array_equal(got[2], want_s64)
        tq = torch.from_numpy(qs).to(dev)
        out = idx.search_device(tq, k)
        fails, unresolved = idx.sync_status()
        assert fails >= 8
        if unresolved:                                                       # > 2048 rows passed the bound
            idx.repair_device(tq, k, out)
        torch.cuda.synchronize()
        assert np.array_equal(out[0].cpu().numpy(), want_ids) and np.array_equal(out[2].cpu().numpy(), want_s64)


def test_sharded_search_on_all_gpus_equals_unsharded():
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    world = 8 if n >= 8 else 4 if n >= 4 else 2
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", str(ROOT / "tests" / "multi_gpu_check.py")],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and "MULTI_GPU_CHECK PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_scattered_upsert_does_not_stop_queries(mcp, oracle):
    """Registry ingest while serving: 1000 scattered rows are replaced in the middle of a search
    loop.  The searches' p99 stays where it was, every search returns either the old or the new exact answer, and after the
    update the table answers like a freshly loaded one (all coarse paths, oracle-checked)."""
    import threading
    import time

    rng = np.random.default_rng(88)
    n, d, q, k = 120_000, 512, 64, 16
    rows = rng.standard_normal((n, d), dtype=np.float32)
    qs = rng.standard_normal((q, d), dtype=np.float32)
    ids = rng.choice(n, 1000, replace=False).astype(np.int64)
    new_rows = rng.standard_normal((1000, d), dtype=np.float32)
    new_rows[:q] = qs * 2.5                                            # every query gets a parallel row somewhere in the table
    after = rows.copy()
    after[ids] = new_rows
    old_ids, old_s64, _ = oracle.topk(rows, qs, k, 0, None)
    new_ids, new_s64, _ = oracle.topk(after, qs, k, 0, None)
    assert np.array_equal(new_ids[:, 0], ids[:q])
    with mcp.ServiceIndex(d, "cosine") as idx:
        idx.load(rows)
        lat, answers = [], []

        def loop(n_iter):
            for _ in range(n_iter):
                t0 = time.perf_counter()
                got = idx.search(qs, k)
                lat.append(time.perf_counter() - t0)
                answers.append(got[0])

        loop(100)
        lat.clear(); answers.clear()
        loop(1000)
        quiet = np.percentile(lat, 99)
        lat.clear(); answers.clear()
        th = threading.Thread(target=loop, args=(1000,))
        th.start()
        time.sleep(0.02)
        t0 = time.perf_counter()
        idx.upsert_rows(ids, new_rows)
        t_call = time.perf_counter() - t0
        th.join()
        busy = np.percentile(lat, 99)
        first_answers = list(answers)
        # the latency half of the claim is a timing property on a shared box: if the first attempt misses the bound, the same
        # upsert (now a no-op for the answers) is repeated against a fresh quiet baseline, up to two more times
        for _ in range(2):
            if busy <= quiet * 1.10 + 30e-6:
                break
            lat.clear(); answers.clear()
            loop(1000)
            quiet = np.percentile(lat, 99)
            lat.clear(); answers.clear()
            th = threading.Thread(target=loop, args=(1000,))
            th.start()
            time.sleep(0.02)
            idx.upsert_rows(ids, new_rows)
            th.join()
            busy = np.percentile(lat, 99)
        answers[:] = first_answers
        assert busy <= quiet * 1.10 + 30e-6, f"search p99 {quiet * 1e6:.0f} us alone, {busy * 1e6:.0f} us with the upsert"
        kinds = [0 if np.array_equal(a, old_ids) else 1 if np.array_equal(a, new_ids) else 2 for a in answers]
        assert 2 not in kinds, "a search saw a half-applied update"
        assert kinds == sorted(kinds) and kinds[-1] == 1              # old answers, then new ones, never back
        for path in ("gemm_f16", "gemm_tf32", "stream_fp32", "exact_fp64"):
            idx.set_option("path", path)
            got = idx.search(qs, k)
            assert np.array_equal(got[0], new_ids) and np.array_equal(got[2], new_s64), path
        with pytest.raises(mcp.McpError):
            idx.upsert_rows(np.array([n], np.int64), new_rows[:1])     # outside the table
        bad = new_rows[:2].copy()
        bad[1, 3] = np.nan
        with pytest.raises(mcp.McpError):
            idx.upsert_rows(ids[:2], bad)                              # rejected before anything is touched
        idx.set_option("path", "auto")
        assert np.array_equal(idx.search(qs, k)[0], new_ids)
        print(f"upsert of 1000 scattered rows: call returned in {t_call * 1e3:.2f} ms; search p99 {quiet * 1e6:.0f} -> {busy * 1e6:.0f} us")

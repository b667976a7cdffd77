"""Parity of the CUDA retrieval path against the CPU oracle, through the C ABI.

Bar (BASELINE.json north_star): top-k ids bit-exact with ties -> lowest id; scores within 1e-5
relative (they are in fact bit-equal: the rescoring kernel repeats the oracle's fp64 arithmetic).
"""
from pathlib import Path

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).parent / "golden"
REL_TOL = 1e-5
PATHS = ["stream_fp32", "gemm_tf32", "gemm_f16", "exact_fp64"]
PATH_CODE = {"stream_fp32": 1, "gemm_tf32": 2, "exact_fp64": 3, "gemm_f16": 4}
COARSE = ["stream_fp32", "gemm_tf32", "gemm_f16"]


def _assert_parity(got, want_ids, want_s64):
    ids, s32, s64 = got
    assert ids.dtype == np.int32 and np.array_equal(ids, want_ids), \
        f"ids differ at {np.argwhere(ids != want_ids)[:5].tolist()}"
    fin = np.isfinite(want_s64)
    assert np.array_equal(np.isnan(s64), np.isnan(want_s64))
    assert np.allclose(s64[fin], want_s64[fin], rtol=REL_TOL, atol=0)
    assert np.array_equal(s64, want_s64, equal_nan=True), "fp64 scores are expected to be bit-equal"
    assert np.array_equal(s32, want_s64.astype(np.float32), equal_nan=True)


def _run(mcp, c, path, k=None, kprime=None):
    with mcp.ServiceIndex(c["rows"].shape[1], "cosine" if c["metric"] == 0 else "l2") as idx:
        idx.load(c["rows"])
        if c["penalty"] is not None:
            idx.set_penalty(c["penalty"])
        idx.set_option("path", path)
        if kprime:
            idx.set_option("kprime", kprime)
        out = idx.search(c["queries"], k or c["k"])
        return out, idx.stats()


@pytest.mark.parametrize("path", PATHS)
@pytest.mark.parametrize("name", ["config1", "small_l2", "edge_cases", "config2_small"])
def test_golden_and_oracle(mcp, oracle, name, path):
    c = {"config1": cases.config1, "small_l2": lambda: cases.small_l2(oracle.telemetry_penalty),
         "edge_cases": lambda: cases.edge_cases(oracle.telemetry_penalty),
         "config2_small": lambda: cases.config2_small(oracle.telemetry_penalty)}[name]()
    k = min(c["k"], 256)
    g = np.load(GOLD / f"retrieval_{name}.npz")
    got, stats = _run(mcp, c, path, k=k)
    _assert_parity(got, g["ids"][:, :k], g["scores64"][:, :k])
    assert stats["last_path"] == PATH_CODE[path]


@pytest.mark.parametrize("metric", [0, 1])
@pytest.mark.parametrize("path", PATHS)
def test_seeded_batches(mcp, oracle, metric, path):
    """ragged sizes: N not a multiple of any tile, Q = 1 .. 37, k from 1 to 256, un-normalised rows."""
    rng = np.random.default_rng(100 + metric)
    for n, d, q, k in [(1, 64, 1, 1), (37, 128, 3, 64), (5000, 256, 37, 100), (4097, 1536, 9, 256), (20000, 768, 16, 10)]:
        rows = (rng.standard_normal((n, d), dtype=np.float32) * rng.uniform(0.5, 2.0, (n, 1)).astype(np.float32))
        qs = rng.standard_normal((q, d), dtype=np.float32)
        pen = (rng.random(n) * 0.05).astype(np.float32)
        c = dict(rows=rows, queries=qs, k=k, metric=metric, penalty=pen)
        want_ids, want_s64, _ = oracle.topk(rows, qs, k, metric, pen)
        got, _ = _run(mcp, c, path)
        _assert_parity(got, want_ids, want_s64)


@pytest.mark.parametrize("path", COARSE)
def test_certificate_catches_near_ties(mcp, oracle, path):
    """Rows that differ from each other by ~1 ulp in score: the coarse pass cannot order them, the
    certificate must notice when k' is too small and the fp64 scan must fix it."""
    rng = np.random.default_rng(7)
    d = 256
    base = rng.standard_normal(d).astype(np.float32)
    rows = np.tile(base, (3000, 1))
    rows += (rng.standard_normal((3000, d)) * 1e-7).astype(np.float32) * np.abs(base)   # near-duplicates
    qs = (base + rng.standard_normal((5, d)).astype(np.float32) * 0.01).astype(np.float32)
    for metric in (0, 1):
        c = dict(rows=rows, queries=qs, k=20, metric=metric, penalty=None)
        want_ids, want_s64, _ = oracle.topk(rows, qs, 20, metric, None)
        got, stats = _run(mcp, c, path, kprime=32)
        _assert_parity(got, want_ids, want_s64)
        assert stats["certificate_failures"] > 0, "adversarial near-ties should have failed the certificate"
        assert stats["exact_scans"] >= stats["certificate_failures"]


def test_zero_query_and_degenerate_inputs(mcp, oracle):
    rng = np.random.default_rng(9)
    rows = rng.standard_normal((500, 64), dtype=np.float32)
    qs = rng.standard_normal((3, 64), dtype=np.float32)
    qs[1] = 0.0                                                   # cosine: every score NaN -> ids 0..k-1
    want_ids, want_s64, _ = oracle.topk(rows, qs, 7, 0, None)
    got, _ = _run(mcp, dict(rows=rows, queries=qs, k=7, metric=0, penalty=None), "auto")
    _assert_parity(got, want_ids, want_s64)
    assert got[0][1].tolist() == list(range(7)) and np.isnan(got[2][1]).all()
    with mcp.ServiceIndex(64) as idx:
        bad = rows.copy()
        bad[3, 5] = np.nan
        with pytest.raises(mcp.McpError):
            idx.load(bad)                                         # pgvector rejects NaN at input; so do we
        idx.load(rows)
        with pytest.raises(mcp.McpError):
            idx.search(qs, 0)
        with pytest.raises(mcp.McpError):
            idx.search(qs, 257)
        with pytest.raises(ValueError):
            idx.search(qs[:, :10], 3)
        ids, _, _ = idx.search(qs[:1], 5)
        idx.set_penalty(np.where(np.arange(500) == ids[0, 0], 100.0, 0.0).astype(np.float32))
        ids2, _, _ = idx.search(qs[:1], 5)
        assert ids2[0, 0] == ids[0, 1]                            # the penalised best service drops out
        idx.set_penalty(None)
        assert np.array_equal(idx.search(qs[:1], 5)[0], ids)


def test_k_exceeds_rows_and_id_base(mcp, oracle):
    rows = np.eye(3, 8, dtype=np.float32)
    qs = np.ones((2, 8), np.float32)
    with mcp.ServiceIndex(8, "l2") as idx:
        idx.load(rows, id_base=1000)
        ids, s32, s64 = idx.search(qs, 6)
    want_ids, want_s64, _ = oracle.topk(rows, qs, 6, 1, None, id_base=1000)
    _assert_parity((ids, s32, s64), want_ids, want_s64)
    assert ids[0].tolist() == [1000, 1001, 1002, -1, -1, -1]


def test_device_pointer_api_and_shard_merge(mcp, oracle):
    """search_device on torch tensors; 4 row shards merged by the K5 kernel == unsharded answer."""
    import torch

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import sharded

    c = cases.config2_small(oracle.telemetry_penalty, n=6000, q=20, k=32)
    want_ids, want_s64, _ = oracle.topk(c["rows"], c["queries"], 32, 0, c["penalty"])
    dev = torch.device("cuda:0")
    rows = torch.from_numpy(c["rows"]).to(dev)
    qs = torch.from_numpy(c["queries"]).to(dev)
    pen = torch.from_numpy(c["penalty"]).to(dev)
    parts_s, parts_i, keep = [], [], []
    for r in range(4):
        lo, hi = sharded.shard_bounds(6000, 4, r)
        idx = mcp.ServiceIndex(1536).adopt(rows[lo:hi], id_base=lo)
        idx.set_penalty(pen[lo:hi].contiguous())
        ids, s32, s64 = idx.search_device(qs, 32)
        assert idx.sync_status()[1] == 0
        parts_s.append(s64)
        parts_i.append(ids)
        keep.append(idx)
    mi, ms32, ms64 = mcp.merge_topk_device(torch.stack(parts_s).contiguous(), torch.stack(parts_i).contiguous())
    torch.cuda.synchronize()
    _assert_parity((mi.cpu().numpy(), ms32.cpu().numpy(), ms64.cpu().numpy()), want_ids, want_s64)
    for idx in keep:
        idx.close()


@pytest.mark.parametrize("path", COARSE)
def test_config2_full_size_properties(mcp, oracle, path):
    """BASELINE config 2 at full size (100K x 1536, 256 queries, top-32, cosine + telemetry).
    The oracle checks a sample of queries; every query is checked against the GPU fp64 scan
    (itself oracle-checked above), plus sortedness and idempotence."""
    import torch

    dev = torch.device("cuda:0")
    n, q, k = 100_000, 256, 32
    g = torch.Generator(device=dev).manual_seed(1002)
    rows = torch.randn((n, 1536), generator=g, device=dev)
    rows /= rows.norm(dim=1, keepdim=True)
    qs = torch.randn((q, 1536), generator=g, device=dev)
    qs /= qs.norm(dim=1, keepdim=True)
    pen_np = oracle.telemetry_penalty(*cases.telemetry(n, 3002))
    idx = mcp.ServiceIndex(1536).adopt(rows)
    idx.set_penalty(torch.from_numpy(pen_np).to(dev))
    idx.set_option("path", path)
    ids, s32, s64 = (t.clone() for t in idx.search_device(qs, k))
    fails, unresolved = idx.sync_status()
    assert unresolved == 0
    ids_b, _, s64_b = idx.search_device(qs, k)
    torch.cuda.synchronize()
    assert torch.equal(ids, ids_b) and torch.equal(s64, s64_b)                   # idempotent
    assert bool((s64[:, 1:] >= s64[:, :-1]).all())                              # ascending
    idx.set_option("path", "exact_fp64")
    ids_x, _, s64_x = idx.search_device(qs, k)
    torch.cuda.synchronize()
    assert torch.equal(ids, ids_x) and torch.equal(s64, s64_x)                   # == brute-force fp64 scan on the GPU
    rows_np, qs_np = rows.cpu().numpy(), qs.cpu().numpy()
    sample = [0, 1, 127, 128, 255]
    want_ids, want_s64, _ = oracle.topk(rows_np, qs_np[sample], k, 0, pen_np)
    _assert_parity((ids.cpu().numpy()[sample], s32.cpu().numpy()[sample], s64.cpu().numpy()[sample]), want_ids, want_s64)
    assert fails <= 2, f"{fails} certificate failures on random data: oversampling is too tight"
    idx.close()


def _rho(v, fmt):
    """Upper bound of |v - round(v)|_2 / |v|_2 per row, as prep.cuh computes it (TableConsts::rho_*, rho_q): tf32 drops the
    low 13 mantissa bits; the fp16 image is the row scaled by the power of two that puts its largest element in [2^14, 2^15),
    rounded to nearest even."""
    v = np.ascontiguousarray(v, np.float32)
    if fmt == "tf32":
        r = (v.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32).astype(np.float64)
    else:
        vmax = np.abs(v).max(axis=1, keepdims=True)
        sc = np.exp2(14.0 - np.floor(np.log2(vmax))).astype(np.float32)
        r = (v * sc).astype(np.float16).astype(np.float64) / sc.astype(np.float64)
    delta = v.astype(np.float64) - r
    n2 = (v.astype(np.float64) ** 2).sum(axis=1)
    return np.sqrt((delta ** 2).sum(axis=1) * (1.0 + 2.0 ** -10) / n2) * (1.0 + 1e-6)


@pytest.mark.parametrize("metric", [0, 1])
@pytest.mark.parametrize("path", COARSE)
def test_coarse_scores_within_certified_bounds(mcp, oracle, metric, path):
    """Every coarse score must lie within the error bound the certificate assumes for its kernel
    (rescore.cuh / mcp_abi.cu CertParams); also localises a wrong TMA / UMMA descriptor to the element.
    Ragged on purpose: N not a multiple of the 256-row tile, Q = 150 (second query block half empty),
    dim = 200 is not a multiple of the 128-byte k-block (TMA zero fill), dim = 1530 is not a multiple of 4 (padded pitch).
    Tensor-core scans: the bound is rho_q + (1 + rho_q) rho_e relative to |q||e|, from the operands' actual rounding
    residuals; on this data it is ~2.4x below the worst-case per-element bound (2^-9 for tf32), and the fp16 shadow's is
    half of tf32's again."""
    import torch

    dev = torch.device("cuda:0")
    u = 2.0 ** -24
    for n, d, q in [(3000, 1536, 150), (777, 200, 5), (600, 1530, 9)]:
        if path == "stream_fp32" and d % 4:
            continue                                   # the validation entry point runs K1 on unpadded queries only
        rng = np.random.default_rng(n + metric)
        rows = (rng.standard_normal((n, d), dtype=np.float32) * rng.uniform(0.5, 2.0, (n, 1)).astype(np.float32))
        qs = rng.standard_normal((q, d), dtype=np.float32)
        pen = (rng.random(n) * 0.05).astype(np.float32)
        exact = np.stack([oracle.scores(rows, qq, metric, pen) for qq in qs])
        with mcp.ServiceIndex(d, "cosine" if metric == 0 else "l2") as idx:
            idx.load(rows).set_penalty(pen)
            got = idx.coarse_scores_device(torch.from_numpy(qs).to(dev), path)
            torch.cuda.synchronize()
            got = got.cpu().numpy().astype(np.float64)
        if path == "stream_fp32":
            rel = 2.0 * (d / 32.0 + 32.0) * u
        else:
            fmt = "tf32" if path == "gemm_tf32" else "f16"
            rq, re = _rho(qs, fmt)[:, None], _rho(rows, fmt).max()
            rel = d * 4.0 * u + 128.0 * u + 1.001 * (rq + (1.0 + rq) * re)
            assert rel.max() < 0.6 * 2.0 ** -9 + d * 4.0 * u, "residual bound is not tighter than the worst case"
        qn = np.linalg.norm(qs.astype(np.float64), axis=1)[:, None]
        en = np.linalg.norm(rows.astype(np.float64), axis=1)[None, :]
        if metric == 0:
            eps = rel + 8 * u * (2.0 + 0.05)
        elif path == "stream_fp32":
            eps = rel * (qn + en.max()) + 8 * u * (qn + en.max() + 0.05)
        else:
            delta = 2.0 * rel * qn * en.max() + 4 * u * (qn + en.max()) ** 2
            d2 = (exact - pen[None, :].astype(np.float64)) ** 2
            eps = np.sqrt(d2 + delta) - np.sqrt(d2) + 8 * u * (np.sqrt(d2) + 0.05) + 1e-6
        err = np.abs(got - exact)
        worst = np.unravel_index(np.argmax(err - eps), err.shape)
        assert (err <= eps).all(), f"{path} metric {metric} n={n} d={d}: |coarse-exact|={err[worst]:.3e} at (q,row)={worst}, bound {np.broadcast_to(eps, err.shape)[worst]:.3e}"


@pytest.mark.parametrize("metric", [0, 1])
def test_gemm_list_prune_path(mcp, oracle, metric):
    """Without the cross-CTA bootstrap every (CTA, query) list starts empty and fills up, so the register-only warp
    radix select that cuts a full list back to k' runs many times; with it, almost never.  Both must give the oracle's
    answer -- this is the only test that drives the prune path hard (one query block, 30K rows per... list)."""
    rng = np.random.default_rng(31 + metric)
    n, d, q, k = 60_000, 128, 40, 24
    rows = rng.standard_normal((n, d), dtype=np.float32)
    qs = rng.standard_normal((q, d), dtype=np.float32)
    pen = (rng.random(n) * 0.02).astype(np.float32)
    want_ids, want_s64, _ = oracle.topk(rows, qs, k, metric, pen)
    for boot in (0, 1):
        with mcp.ServiceIndex(d, "cosine" if metric == 0 else "l2") as idx:
            idx.load(rows).set_penalty(pen)
            idx.set_option("path", "gemm_tf32").set_option("gemm_bootstrap", boot).set_option("kprime", 32)
            _assert_parity(idx.search(qs, k), want_ids, want_s64)
            assert idx.stats()["last_kprime"] == 32


def test_micro_batched_planner_matches_direct_search(mcp, oracle):
    import threading

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    rows = cases.unit_rows(20_000, 1536, 77)
    names = [f"s{i}" for i in range(len(rows))]

    class Reg:
        def list_services(self):
            return []

    planner = cp.GraphPlanner(Reg(), fetch_embeddings=lambda: list(zip(names, rows)), top_k=6, micro_batch=True)
    planner.refresh_index([])
    intents = [f"intent {i}" for i in range(96)]
    got = [None] * len(intents)

    def client(i):
        got[i] = planner.retrieve(intents[i])

    th = [threading.Thread(target=client, args=(i,)) for i in range(len(intents))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    qs = np.stack([cp.hash_embedding(t) for t in intents])
    want_ids, want_s64, _ = oracle.topk(rows, qs, 6, 0, np.zeros(len(rows), np.float32))
    assert [[n for n, _ in r] for r in got] == [[names[i] for i in row] for row in want_ids]
    assert planner._batcher.requests == 96 and planner._batcher.launches <= 96


def test_update_rows_and_table_file(mcp, oracle, tmp_path):
    """Registry refresh without a reload: replaced rows are found (or no longer found) at once, on a loaded table
    and on an adopted one; and a table written to the flat file format searches the same after load_file."""
    import torch

    rng = np.random.default_rng(55)
    rows = rng.standard_normal((5000, 256), dtype=np.float32)
    qs = rng.standard_normal((12, 256), dtype=np.float32)
    with mcp.ServiceIndex(256) as idx:
        idx.load(rows)
        before, _, _ = idx.search(qs, 4)
        new = rows.copy()
        new[1000:1012] = qs * 3.0                                      # each query now has a parallel row
        new[int(before[0, 0])] = -qs[0]                                # and query 0's old best becomes its worst
        lo, hi = min(1000, int(before[0, 0])), max(1012, int(before[0, 0]) + 1)
        idx.update_rows(lo, new[lo:hi])
        want_ids, want_s64, _ = oracle.topk(new, qs, 4, 0, None)
        _assert_parity(idx.search(qs, 4), want_ids, want_s64)
        assert want_ids[:, 0].tolist() == list(range(1000, 1012))
        bad = new[:3].copy()
        bad[1, 1] = np.inf
        with pytest.raises(mcp.McpError):
            idx.update_rows(0, bad)
        idx.update_rows(0, new[:3])
    dev = torch.device("cuda:0")
    t = torch.from_numpy(rows).to(dev)
    with mcp.ServiceIndex(256) as idx:
        idx.adopt(t)
        t[1000:1012] = torch.from_numpy(qs * 3.0).to(dev)
        torch.cuda.synchronize()
        idx.update_rows(1000, None, 12)
        got = idx.search(qs, 1)[0]
        assert got[:, 0].tolist() == list(range(1000, 1012))
        with pytest.raises(mcp.McpError):
            idx.update_rows(0, rows[:2])                               # adopted memory belongs to the caller
    p = tmp_path / "t.mcp"
    mcp.ServiceIndex.save_table(p, new, [f"n{i}" for i in range(len(new))])
    with mcp.ServiceIndex(256) as idx:
        names = idx.load_file(p)
        assert names[7] == "n7"
        _assert_parity(idx.search(qs, 4), want_ids, want_s64)
        # a row window of the file (what one rank of a sharded registry reads): ids stay the file's row numbers
        names = idx.load_file(p, 900, 2200)
        assert idx.n_rows == 2200 and names[0] == "n900" and len(names) == 2200
        w_ids, w_s64, _ = oracle.topk(new[900:3100], qs, 4, 0, None, id_base=900)
        _assert_parity(idx.search(qs, 4), w_ids, w_s64)
        with pytest.raises(mcp.McpError):
            idx.load_file(p, 4000, 2000)                               # past the end of the file
    odd = rng.standard_normal((777, 250), dtype=np.float32)            # dim % 4 != 0: rows are padded on the way in
    p2 = tmp_path / "odd.mcp"
    mcp.ServiceIndex.save_table(p2, odd, id_base=10)
    q2 = rng.standard_normal((3, 250), dtype=np.float32)
    o_ids, o_s64, _ = oracle.topk(odd, q2, 5, 1, None, id_base=10)
    with mcp.ServiceIndex(250, "l2") as idx:
        idx.load_file(p2)
        _assert_parity(idx.search(q2, 5), o_ids, o_s64)
    with mcp.ServiceIndex(256, "cosine", [0, 0]) as multi:             # two shards, each reading its own half of the file
        multi.load_file(p)
        _assert_parity(multi.search(qs, 4), want_ids, want_s64)


def test_query_chunking_paths_agree(mcp, oracle):
    """Batches larger than one launch takes: the GEMM path splits at 1024 queries, the streaming path at whatever keeps
    its score matrix within budget.  Every chunked answer must equal the fp64 scan's (oracle-checked on a sample)."""
    import torch

    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(99)
    n, d, q, k = 150_000, 512, 1300, 16
    rows = torch.randn((n, d), generator=g, device=dev)
    qs = torch.randn((q, d), generator=g, device=dev)
    idx = mcp.ServiceIndex(d).adopt(rows)
    idx.set_option("path", "exact_fp64")
    want_i, _, want_s = (t.clone() for t in idx.search_device(qs, k))
    idx.sync_status()
    for path in ("gemm_tf32", "gemm_f16", "stream_fp32"):
        idx.set_option("path", path)
        ids, _, s64 = idx.search_device(qs, k)
        assert idx.sync_status()[1] == 0
        assert torch.equal(ids, want_i) and torch.equal(s64, want_s), path
    sample = [0, 1023, 1024, 1299]
    oi, os64, _ = oracle.topk(rows.cpu().numpy(), qs.cpu().numpy()[sample], k, 0, None)
    assert np.array_equal(want_i.cpu().numpy()[sample], oi) and np.array_equal(want_s.cpu().numpy()[sample], os64)
    idx.close()


@pytest.mark.parametrize("path", ["gemm_tf32", "gemm_f16"])
@pytest.mark.parametrize("metric", [0, 1])
def test_cta_pair_and_single_cta_scans_agree_with_oracle(mcp, oracle, metric, path):
    """K2 has three shapes: one tcgen05 CTA pair per cluster (cta_group::2, the default for an even number of 128-query
    blocks), two single CTAs multicasting the row tile, and plain single CTAs.  Same oracle answer from each, with and
    without balanced tile heights (which change the MMA's N and the TMA box)."""
    rng = np.random.default_rng(101 + metric)
    n, d, q, k = 70_001, 96, 256, 16           # > 65536 rows: the bootstrap tile hands its accumulator back early
    rows = rng.standard_normal((n, d), dtype=np.float32)
    qs = rng.standard_normal((q, d), dtype=np.float32)
    pen = (rng.random(n) * 0.02).astype(np.float32)
    want_ids, want_s64, _ = oracle.topk(rows, qs, k, metric, pen)
    for opts in ({"gemm_pair": 1}, {"gemm_pair": 0}, {"gemm_cluster": 1}, {"gemm_pair": 1, "gemm_balance": 1}):
        with mcp.ServiceIndex(d, "cosine" if metric == 0 else "l2") as idx:
            idx.load(rows).set_penalty(pen)
            idx.set_option("path", path)
            for name, v in opts.items():
                idx.set_option(name, v)
            _assert_parity(idx.search(qs, k), want_ids, want_s64)
            assert idx.stats()["last_path"] == PATH_CODE[path], opts


def test_bootstrap_tile_drops_are_caught_by_the_certificate(mcp, oracle):
    """On large tables K2's first tile keeps only four rows per (query, half tile) and records the best score it may
    have dropped; the certificate must then refuse any answer such a row could belong to.  Here the whole true top-k
    of query 0 sits inside one half of the first tile (rows 0..39 are near-copies of the query), so 36 of them are
    dropped by construction: the query has to fail the certificate and come back exact from the fp64 scan."""
    rng = np.random.default_rng(202)
    n, d, q, k = 70_000, 64, 256, 32
    rows = rng.standard_normal((n, d), dtype=np.float32)
    qs = rng.standard_normal((q, d), dtype=np.float32)
    rows[:40] = qs[0] + rng.standard_normal((40, d)).astype(np.float32) * 0.05
    rows[300:330] = qs[130] + rng.standard_normal((30, d)).astype(np.float32) * 0.05   # second query block, next CTA's first tile
    want_ids, want_s64, _ = oracle.topk(rows, qs, k, 0, None)
    assert set(range(40)).issuperset(want_ids[0].tolist())
    with mcp.ServiceIndex(d, "cosine") as idx:
        idx.load(rows)
        idx.set_option("path", "gemm_tf32")
        _assert_parity(idx.search(qs, k), want_ids, want_s64)
        st = idx.stats()
        assert st["last_path"] == 2
        assert st["certificate_failures"] >= 1, "the dropped rows of the bootstrap tile went unnoticed"


def test_foreign_memory_is_refused(mcp):
    """Every device pointer handed to an index must be plain device memory of the index's own GPU (the kernels and the TMA
    tensor maps built over those buffers read this GPU's memory only): pinned host memory, and on a multi-GPU box another
    GPU's memory, are refused with MCP_EINVAL at the C ABI -- the Python wrapper's own checks are bypassed here."""
    import ctypes as C

    import torch

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import _native as nat

    lib = nat.lib()
    d = 64
    dev = torch.device("cuda:0")
    rows = torch.randn((1000, d), device=dev)
    with mcp.ServiceIndex(d, "cosine", 0) as idx:
        idx.adopt(rows)
        ids = torch.empty((4, 3), dtype=torch.int32, device=dev)
        sc = torch.empty((4, 3), dtype=torch.float32, device=dev)
        host_q = torch.randn((4, d)).pin_memory()
        rc = lib.mcp_index_search_device(idx._h, C.c_void_p(host_q.data_ptr()), 4, 3, C.c_void_p(ids.data_ptr()), C.c_void_p(sc.data_ptr()), None, None)
        assert rc == -1 and b"not device memory of GPU 0" in lib.mcp_last_error()
        host_rows = torch.randn((100, d)).pin_memory()
        rc = lib.mcp_index_adopt_device(idx._h, C.c_void_p(host_rows.data_ptr()), 100, d, 0, None)
        assert rc == -1
        if torch.cuda.device_count() >= 2:
            peer = torch.randn((1000, d), device="cuda:1")
            rc = lib.mcp_index_adopt_device(idx._h, C.c_void_p(peer.data_ptr()), 1000, d, 0, None)
            assert rc == -1 and b"device 1" in lib.mcp_last_error()
            peer_q = torch.randn((4, d), device="cuda:1")
            rc = lib.mcp_index_search_device(idx._h, C.c_void_p(peer_q.data_ptr()), 4, 3, C.c_void_p(ids.data_ptr()), C.c_void_p(sc.data_ptr()), None, None)
            assert rc == -1
        q = torch.randn((4, d), device=dev)                                  # and the index still works afterwards
        got = idx.search_device(q, 3)[0]
        torch.cuda.synchronize()
        assert got.shape == (4, 3) and int(got.min()) >= 0


@pytest.mark.parametrize("metric", [0, 1])
def test_gpu_answer_within_tolerance_of_pgvector_style_fp32(mcp, oracle, metric):
    """The bar as north_star words it: against an exact fp32 brute-force reference of pgvector's semantics (pgvector's own float
    accumulators, oracle.c), scores within 1e-5 relative and ids equal except between rows that reference cannot tell apart."""
    rng = np.random.default_rng(610 + metric)
    rows = rng.standard_normal((30_000, 1536), dtype=np.float32)
    rows /= np.linalg.norm(rows, axis=1, keepdims=True)
    qs = rng.standard_normal((16, 1536), dtype=np.float32)
    pen = (rng.random(30_000) * 0.02).astype(np.float32)
    with mcp.ServiceIndex(1536, "cosine" if metric == 0 else "l2") as idx:
        idx.load(rows).set_penalty(pen)
        ids, s32, s64 = idx.search(qs, 32)
    for i in range(qs.shape[0]):
        ref = oracle.pgvector_f32_scores(rows, qs[i], metric, pen)
        assert oracle.agrees_with_fp32_reference(ids[i], s64[i], ref, 32), i

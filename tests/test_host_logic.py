"""Host-side logic and the C-ABI surface, no GPU needed."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _declared_symbols():
    text = (ROOT / "include" / "mcp_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mcp_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol(mcp):
    nat = mcp._native
    lib = nat.lib()                       # builds with nvcc if stale; raises loudly if it cannot
    declared = _declared_symbols()
    assert len(declared) >= 18
    raw = ctypes.CDLL(str(nat.LIB_PATH))
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in include/mcp_b200.h but not exported"
        assert name in nat.SIGNATURES, f"{name} has no ctypes signature"
    assert set(nat.SIGNATURES) == set(declared)
    assert b"sm_100a" in lib.mcp_build_info()


def test_sass_is_sm100a_only(mcp):
    import shutil
    import subprocess

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    out = subprocess.run([cuobjdump, "-lelf", str(mcp._native.LIB_PATH)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback_without_device(mcp):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert mcp.device_count() == 0
    with pytest.raises(mcp.McpError) as e:
        mcp.ServiceIndex(16)
    assert e.value.code == -2            # MCP_ECUDA: fails loudly, nothing computes on the CPU
    enc = mcp.dag.encode_graphs([{"nodes": [{"name": "a"}], "edges": []}])
    with pytest.raises(mcp.McpError):
        mcp.dag.levels(enc)


def test_argument_validation_without_device(mcp):
    lib = mcp._native.lib()
    assert lib.mcp_index_create(None, 0, 16, 0) == -1
    h = ctypes.c_void_p()
    assert lib.mcp_index_create(ctypes.byref(h), 0, 0, 0) == -1
    assert b"dim" in lib.mcp_last_error()
    assert lib.mcp_index_create(ctypes.byref(h), 0, 16, 7) == -1
    assert lib.mcp_index_search(None, None, 1, 1, None, None, None) == -1
    assert lib.mcp_merge_topk_device(None, None, 1, 1, 1, None, None, None, None) == -1
    assert lib.mcp_dag_levels(0, None, None, None, 1, None, None, None, None) == -1
    assert lib.mcp_dag_levels(0, None, None, None, 0, None, None, None, None) == -1


def test_encode_graphs_follows_digraph_insertion_rules(mcp):
    g = {"nodes": [{"name": "b"}, {"name": "a"}, {"name": "b"}],
         "edges": [{"from": "x", "to": "a"}, {"from": "a", "to": "y"}, {"from": "x", "to": "a"}]}
    enc = mcp.dag.encode_graphs([g, {"nodes": [], "edges": []}])
    assert enc.names[0] == ["b", "a", "x", "y"]              # declared first (dupe keeps position), then u, v as seen
    assert enc.n_nodes.tolist() == [4, 0]
    assert enc.edge_off.tolist() == [0, 3, 3]
    assert enc.edges.tolist() == [[2, 1], [1, 3], [2, 1]]     # the duplicate edge stays; the kernel collapses it
    assert mcp.dag.encode_graphs([{"nodes": [{"name": f"n{i}"} for i in range(65)], "edges": []}]).n_nodes.tolist() == [65]   # the wide kernel's
    big = {"nodes": [{"name": f"n{i}"} for i in range(256)], "edges": []}
    with pytest.raises(ValueError):
        mcp.dag.encode_graphs([big])


def test_synthetic_dags_shape(mcp):
    enc = mcp.dag.synthetic_dags(3000, seed=1)
    assert len(enc) == 3000 and enc.edge_off[-1] == enc.edges.shape[0]
    assert 1 <= enc.n_nodes.min() and enc.n_nodes.max() <= 64
    for i in range(50):
        e = enc.edges[enc.edge_off[i]:enc.edge_off[i + 1]]
        assert (e < enc.n_nodes[i]).all()
    mean_e = enc.edges.shape[0] / 3000
    assert 70 < mean_e < 115                                  # about 92 edges per DAG


def test_telemetry_fold_matches_oracle(mcp, oracle):
    import cases

    lat, err, cost = cases.telemetry(5000, 3)
    a = mcp.telemetry.fold_penalty(lat, err, cost)
    b = oracle.telemetry_penalty(lat, err, cost)
    assert a.dtype == np.float32 and np.array_equal(a, b)
    svc = [{"name": "s0", "cost_profile": 0.005}, {"name": "s1"}]

    class R:
        def get(self, k):
            return b'{"latency_ms": 2500, "error_rate": 0.1}' if k.endswith("s0") else None

    p = mcp.telemetry.penalty_for_services(svc, R())
    assert np.allclose(p, [0.05 * 1.0 + 0.5 * 0.1 + 0.005, 0.0])


def test_shard_bounds_and_cpu_merge(mcp, oracle):
    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import sharded

    for n, w in [(10, 3), (1000, 8), (7, 8), (10_000_000, 8)]:
        spans = [sharded.shard_bounds(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    rng = np.random.default_rng(5)
    sc = np.sort(rng.standard_normal((4, 6, 9)), axis=2)
    sc[1, 2, 7:] = np.nan
    ids = rng.permutation(4 * 6 * 9).astype(np.int32).reshape(4, 6, 9)
    sc[3, :, 5:] = np.inf
    ids[3, :, 5:] = -1
    sc[0, 0, 0] = sc[2, 0, 0] = -9.0                         # exact tie across shards -> lowest id
    mi, ms = sharded.merge_gathered_cpu(sc, ids)
    oi, os_ = oracle.merge(sc, ids)
    assert np.array_equal(mi, oi) and np.array_equal(ms, os_, equal_nan=True)
    assert mi[0, 0] == min(ids[0, 0, 0], ids[2, 0, 0])


def test_control_plane_surface(mcp):
    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    assert cp.SERVICES_PREFIX == "mcp:service:"
    assert cp.PlanRequest(intent="x").intent == "x"
    assert set(cp.ExecuteResponse(results={}, errors={}).model_dump()) == {"results", "errors"}
    v = cp.hash_embedding("book a flight")
    assert v.shape == (1536,) and v.dtype == np.float32 and np.array_equal(v, cp.hash_embedding("book a flight"))

    class FakeRedis:
        def __init__(self, d):
            self.d = d

        def scan_iter(self, pattern):
            return [k for k in self.d if k.startswith(pattern.rstrip("*"))]

        def get(self, k):
            return self.d.get(k)

    reg = cp.ServiceRegistry(client=FakeRedis({"mcp:service:a": '{"name": "a", "endpoint": "http://a"}', "other": "1"}))
    assert reg.list_services() == [{"name": "a", "endpoint": "http://a"}]


def test_micro_batcher_coalesces_and_answers_each_caller(mcp):
    """Host logic only (a numpy stand-in for the index): concurrent single queries share launches, every caller gets
    the answer to *its* query, a shorter k is a prefix, and an error reaches every waiting caller."""
    import threading
    import time

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    table = np.random.default_rng(0).standard_normal((500, 16)).astype(np.float32)
    calls = []

    def fake_search(q, k):
        calls.append(q.shape[0])
        time.sleep(0.01)                                   # a launch in flight: later requests pile up behind it
        d = ((q[:, None, :] - table[None]) ** 2).sum(-1)
        ids = np.argsort(d, axis=1, kind="stable")[:, :k].astype(np.int32)
        return ids, np.take_along_axis(d, ids, 1).astype(np.float32), None

    mb = cp.MicroBatcher(fake_search, max_batch=64)
    qs = np.random.default_rng(1).standard_normal((200, 16)).astype(np.float32)
    out = [None] * 200

    def client(i):
        out[i] = mb.search(qs[i], 3 + i % 4)

    th = [threading.Thread(target=client, args=(i,)) for i in range(200)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    for i in range(200):
        want = np.argsort(((qs[i] - table) ** 2).sum(-1), kind="stable")[:3 + i % 4]
        assert out[i][0].tolist() == want.tolist()
    assert sum(calls) == 200 and len(calls) < 60 and max(calls) <= 64     # coalesced, capped
    assert mb.requests == 200 and mb.launches == len(calls)
    mb.close()
    boom = cp.MicroBatcher(lambda q, k: (_ for _ in ()).throw(ValueError("no device")))
    with pytest.raises(ValueError):
        boom.search(qs[0], 3)
    boom.close()


def test_table_file_format_round_trip(mcp, tmp_path):
    rows = np.random.default_rng(3).standard_normal((37, 24)).astype(np.float32)
    names = [f"svc-{i}" for i in range(37)]
    p = tmp_path / "table.mcp"
    mcp.ServiceIndex.save_table(p, rows, names, id_base=100, metric=1)
    assert p.stat().st_size == 64 + rows.nbytes
    got, got_names, id_base, metric = mcp.ServiceIndex.open_table(p)
    assert isinstance(got, np.memmap) and np.array_equal(got, rows)
    assert got_names == names and id_base == 100 and metric == 1
    (tmp_path / "junk").write_bytes(b"x" * 100)
    with pytest.raises(ValueError):
        mcp.ServiceIndex.open_table(tmp_path / "junk")


def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` runs on the host cores only (here, no GPU): one JSON line with the contract's keys."""
    import json
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--rows", "3000", "--queries", "16",
                        "--topk", "8", "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "top-k retrieval queries/sec" and d["unit"] == "queries/s"
    for key in ("value", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
                "data", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["value"] > 0

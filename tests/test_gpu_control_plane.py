"""The reference-facing surface on the new path: planner retrieval with a stubbed LLM / registry /
embedding table, orchestrator execution with mocked HTTP services (BASELINE config 5, scaled down)."""
import asyncio
import json

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


class FakeRedis:
    def __init__(self, d):
        self.d = d

    def scan_iter(self, pattern):
        return [k for k in self.d if k.startswith(pattern.rstrip("*"))]

    def get(self, k):
        return self.d.get(k)


def _world(n=3000, d=1536):
    rows = cases.unit_rows(n, d, 1005)
    names = [f"svc{i:05d}" for i in range(n)]
    lat, err, cost = cases.telemetry(n, 2005)
    store = {}
    for i, nm in enumerate(names):
        store[f"mcp:service:{nm}"] = json.dumps({"name": nm, "endpoint": f"http://{nm}/api", "input_schema": {"x": "str"},
                                                 "output_schema": {"y": "str"}, "cost_profile": float(cost[i]), "fallback": None})
        store[f"mcp:telemetry:{nm}"] = json.dumps({"latency_ms": float(lat[i]), "error_rate": float(err[i])})
    return rows, names, store, (lat, err, cost)


def test_plan_retrieves_topk_and_prompts_with_candidates_only(mcp, oracle):
    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    rows, names, store, (lat, err, cost) = _world()
    redis = FakeRedis(store)
    prompts = []

    def llm(prompt):
        prompts.append(prompt)
        return json.dumps({"nodes": [], "edges": []})

    planner = cp.GraphPlanner(cp.ServiceRegistry(client=redis), fetch_embeddings=lambda: list(zip(names, rows)),
                              llm=llm, top_k=8, telemetry_client=redis)
    resp = planner.plan("translate a document and email it")
    assert resp.graph == {"nodes": [], "edges": []}
    q = cp.hash_embedding("translate a document and email it")[None]
    pen = oracle.telemetry_penalty(lat, err, np.asarray(cost, np.float32))
    want_ids, want_s64, _ = oracle.topk(rows, q, 8, 0, pen)
    listed = [ln.split(" ")[1] for ln in prompts[0].splitlines() if ln.startswith("- ")]
    assert listed == [names[i] for i in want_ids[0]]                 # exactly the k candidates, in rank order
    got = planner.retrieve("translate a document and email it")
    assert [n for n, _ in got] == listed
    assert np.allclose([s for _, s in got], want_s64[0], rtol=1e-5)
    batch = planner.retrieve_batch([f"intent {i}" for i in range(40)], 5)
    qs = np.stack([cp.hash_embedding(f"intent {i}") for i in range(40)])
    bi, _, _ = oracle.topk(rows, qs, 5, 0, pen)
    assert [[n for n, _ in r] for r in batch] == [[names[i] for i in row] for row in bi]


def _orch(mcp, handler):
    import httpx

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    return cp, cp.Orchestrator(client=httpx.AsyncClient(transport=httpx.MockTransport(handler)))


def test_execute_order_matches_networkx_and_endpoints(mcp, oracle):
    import httpx
    from fastapi.testclient import TestClient

    calls = []

    def handler(request):
        calls.append(str(request.url))
        if "bad" in str(request.url):
            return httpx.Response(500)
        return httpx.Response(200, json={"from": str(request.url)})

    cp, orch = _orch(mcp, handler)
    graph = {"nodes": [{"name": n, "endpoint": f"http://{n}/api", "inputs": {"z": "z"} if n != "z" else {}} for n in "zyxwv"],
             "edges": [{"from": a, "to": b} for a, b in [("y", "v"), ("z", "w"), ("z", "x"), ("y", "x"), ("z", "v")]]}
    _, order, _ = oracle.nx_levels(graph)
    app = cp.create_app(registry=object(), planner=type("P", (), {"plan": lambda self, i: cp.PlanResponse(graph=graph)})(), orch=orch)
    client = TestClient(app, raise_server_exceptions=False)
    r = client.post("/execute", json={"graph": graph, "payload": {}})
    assert r.status_code == 200 and r.json()["errors"] == {}
    assert calls == [f"http://{n}/api" for n in order]               # == list(nx.topological_sort(G))
    assert r.json()["results"]["x"] == {"from": "http://x/api"}
    calls.clear()
    r = client.post("/plan_and_execute", json={"intent": "anything"})
    assert r.status_code == 200 and len(calls) == 5
    assert client.post("/execute", json={"graph": graph}).status_code == 422          # missing payload
    # root fails, no in-edge -> 502 as in the reference
    g2 = {"nodes": [{"name": "bad", "endpoint": "http://bad/api", "inputs": {}}], "edges": []}
    r = client.post("/execute", json={"graph": g2, "payload": {}})
    assert r.status_code == 502 and "bad failed and no fallback available" in r.json()["detail"]
    # failing node with an in-edge carrying a fallback: the reference 500s on a KeyError (its :119); here the fallback runs
    g3 = {"nodes": [{"name": "a", "endpoint": "http://a/api", "inputs": {}}, {"name": "bad", "endpoint": "http://bad/api", "inputs": {}}],
          "edges": [{"from": "a", "to": "bad", "fallback": "http://spare/api"}]}
    r = client.post("/execute", json={"graph": g3, "payload": {}})
    assert r.status_code == 200 and r.json()["results"]["bad"] == {"from": "http://spare/api"} and "bad" in r.json()["errors"]
    # cycle: acyclic prefix is POSTed, then NetworkXUnfeasible -> 500
    calls.clear()
    g4 = {"nodes": [{"name": n, "endpoint": f"http://{n}/api", "inputs": {}} for n in "abc"],
          "edges": [{"from": "a", "to": "b"}, {"from": "b", "to": "c"}, {"from": "c", "to": "b"}]}
    assert client.post("/execute", json={"graph": g4, "payload": {}}).status_code == 500
    assert calls == ["http://a/api"]


def test_schedule_batch(mcp, oracle):
    cp, orch = _orch(mcp, lambda request: None)
    enc0 = mcp.dag.synthetic_dags(2000, seed=12)
    graphs = [mcp.dag.to_graph_dict(enc0, i) for i in range(len(enc0))]
    enc, res = orch.schedule_batch(graphs)
    for i in (0, 1, 500, 1999):
        assert mcp.dag.decode(enc, res, i) == oracle.nx_levels(graphs[i])
    asyncio.run(orch.client.aclose())


def test_parallel_levels_and_retries(mcp, oracle):
    """Opt-in extras: a generation's nodes are awaited together (same results, calls grouped by generation), and a
    flaky service is retried before its fallback."""
    import httpx

    calls, fail_once = [], {"http://w/api": 1}

    def handler(request):
        url = str(request.url)
        calls.append(url)
        if fail_once.get(url, 0) > 0:
            fail_once[url] -= 1
            return httpx.Response(503)
        return httpx.Response(200, json={"from": url})

    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    orch = cp.Orchestrator(client=httpx.AsyncClient(transport=httpx.MockTransport(handler)), parallel_levels=True, retries=1)
    graph = {"nodes": [{"name": n, "endpoint": f"http://{n}/api", "inputs": {}} for n in "zyxwv"],
             "edges": [{"from": a, "to": b} for a, b in [("y", "v"), ("z", "w"), ("z", "x"), ("y", "x"), ("z", "v")]]}
    resp = asyncio.run(orch.execute(graph, {}))
    assert resp.errors == {} and set(resp.results) == set("zyxwv")     # w failed once, the retry succeeded
    assert calls.count("http://w/api") == 2
    first_gen = {"http://z/api", "http://y/api"}
    assert set(calls[:2]) == first_gen and not (set(calls[2:]) & first_gen)   # generation 0 completes before generation 1 starts

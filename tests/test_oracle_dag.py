"""DAG-levels oracle pinned against the live NetworkX (the reference's call at control_plane.py:104)
and against the committed NetworkX-generated golden file."""
from pathlib import Path

import numpy as np
from hypothesis import given, settings, strategies as st

GOLD = Path(__file__).parent / "golden"


def _check_against_nx(oracle, dag, graphs):
    enc = dag.encode_graphs(graphs)
    level, order, nlev, status = oracle.dag_levels(enc.n_nodes, enc.edge_off, enc.edges)
    for i, g in enumerate(graphs):
        lv, od, stt = oracle.nx_levels(g)
        names = enc.names[i]
        assert status[i] == stt
        assert {names[v]: int(level[i, v]) for v in range(len(names)) if level[i, v] != 0xFF} == lv
        assert [names[v] for v in order[i] if v != 0xFF] == od
        assert nlev[i] == ((max(lv.values()) + 1) if lv else 0)


def test_golden_file(oracle):
    g = np.load(GOLD / "dag_golden.npz")
    level, order, nlev, status = oracle.dag_levels(g["n_nodes"], g["edge_off"], g["edges"])
    assert np.array_equal(level, g["level"])
    assert np.array_equal(order, g["order"])
    assert np.array_equal(nlev, g["n_levels"])
    assert np.array_equal(status, g["status"])
    assert status.sum() >= 2


def test_survey_example_and_edge_cases(oracle, mcp):
    dag = mcp.dag
    ex = {"nodes": [{"name": n} for n in "zyxwv"],
          "edges": [{"from": a, "to": b} for a, b in [("y", "v"), ("z", "w"), ("z", "x"), ("y", "x"), ("z", "v")]]}
    enc = dag.encode_graphs([ex])
    level, order, nlev, status = oracle.dag_levels(enc.n_nodes, enc.edge_off, enc.edges)
    names = enc.names[0]
    assert [names[v] for v in order[0] if v != 0xFF] == ["z", "y", "w", "v", "x"]      # [[z,y],[w,v,x]]
    assert nlev[0] == 2 and status[0] == 0
    _check_against_nx(oracle, dag, [
        ex,
        {"nodes": [], "edges": []},
        {"nodes": [{"name": "a"}], "edges": [{"from": "a", "to": "a"}]},
        {"nodes": [{"name": "a"}, {"name": "b"}], "edges": [{"from": "a", "to": "b"}] * 3},
        {"nodes": [{"name": "a"}], "edges": [{"from": "q", "to": "a"}, {"from": "a", "to": "r"}]},
        {"nodes": [{"name": "a"}, {"name": "b"}, {"name": "a"}], "edges": [{"from": "b", "to": "a"}]},
    ])


def test_synthetic_batch_against_nx(oracle, mcp):
    enc = mcp.dag.synthetic_dags(400, seed=77)
    _check_against_nx(oracle, mcp.dag, [mcp.dag.to_graph_dict(enc, i) for i in range(len(enc))])


@st.composite
def _graphs(draw):
    n = draw(st.integers(0, 12))
    names = [f"s{i}" for i in range(n + 3)]                      # 3 names never declared: implicit nodes
    declared = [{"name": draw(st.sampled_from(names[:max(n, 1)]))} for _ in range(n)]
    m = draw(st.integers(0, 30))
    edges = [{"from": draw(st.sampled_from(names)), "to": draw(st.sampled_from(names))} for _ in range(m)]
    return {"nodes": declared, "edges": edges}


@settings(max_examples=150, deadline=None)
@given(_graphs())
def test_random_graphs_against_nx(g):
    import oracle
    import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as mcp

    _check_against_nx(oracle, mcp.dag, [g])

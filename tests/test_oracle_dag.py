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


def _fifo_restatement(n, edges):
    """What kahn_fifo_kernel's thread does for one DAG (csrc/kahn.cuh), step for step, in Python: bucket the valid edges by
    source in edge order, flag repeats within a source's list, count distinct parents, then a FIFO.  The queue is the order,
    a node's level its releasing parent's + 1."""
    start = [0] * 64
    for u, v in edges:
        if u < n and v < n:
            start[u] += 1
    acc = 0
    for u in range(n):
        c, start[u] = start[u], acc
        acc += c
    pool = [0] * (acc + 1)
    for u, v in edges:
        if u < n and v < n:
            pool[start[u]] = v
            start[u] += 1                                   # ends as the END of u's list
    beg = 0
    for u in range(n):
        if start[u] > beg:
            pool[beg] |= 0x40
        beg = start[u]
    indeg, seen = [0] * 64, 0
    for r in range(acc):
        c = pool[r]
        if c & 0x40:
            seen = 0
        bit = 1 << (c & 0x3F)
        if seen & bit:
            pool[r] = c | 0x80
        else:
            indeg[c & 0x3F] += 1
        seen |= bit
    order, level, tail = [0xFF] * 64, [0xFF] * 64, 0
    for v in range(n):
        if indeg[v] == 0:
            order[tail], level[v], tail = v, 0, tail + 1
    head, lu, bound, nlev, r, rend = 0, 0, tail, 0, 0, 0
    while True:
        if r == rend:
            if head == tail:
                break
            if head == bound:
                lu, bound = lu + 1, tail
            u, head, nlev = order[head], head + 1, lu + 1
            r, rend = (start[u - 1] if u else 0), start[u]
        if r < rend:
            c, r = pool[r], r + 1
            if not c & 0x80:
                v = c & 0x3F
                indeg[v] -= 1
                if indeg[v] == 0:
                    order[tail], level[v], tail = v, lu + 1, tail + 1
    return level, order, nlev, int(tail != n)


def test_the_thread_per_dag_algorithm_is_the_oracles(oracle, mcp):
    """The large-batch kernel does not compute generations from predecessor masks as the oracle does; its algorithm, restated
    here, gives the same bytes on DAGs with repeated edges, self loops, cycles and out-of-range edges."""
    base = mcp.dag.synthetic_dags(1500, seed=7)
    rng = np.random.default_rng(1)
    parts, off = [], [0]
    for i in range(len(base.n_nodes)):
        e = base.edges[base.edge_off[i]:base.edge_off[i + 1]]
        if i % 3 == 0 and len(e):
            e = np.concatenate([e, e[rng.permutation(len(e))[:max(1, len(e) // 2)]]])
            e = e[rng.permutation(len(e))]
        if i % 50 == 0:
            e = np.concatenate([e, np.array([[0, 0]], np.uint8)])
        if i % 70 == 0:
            e = np.concatenate([e, np.array([[200, 1], [1, 99]], np.uint8)])
        parts.append(e)
        off.append(off[-1] + len(e))
    edges, off = np.concatenate(parts), np.asarray(off, np.uint32)
    level, order, nlev, status = oracle.dag_levels(base.n_nodes, off, edges)
    assert status.sum() > 10
    for i in range(len(base.n_nodes)):
        got = _fifo_restatement(int(base.n_nodes[i]), [tuple(x) for x in edges[off[i]:off[i + 1]].tolist()])
        assert got == (list(level[i]), list(order[i]), int(nlev[i]), int(status[i])), i

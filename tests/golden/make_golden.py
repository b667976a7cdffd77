"""Regenerates tests/golden/*.npz.  Run from the repo root: python tests/golden/make_golden.py

retrieval_*.npz  expected (ids, scores64) from the CPU oracle on seeded inputs.  The reference
                 holds no vectors for this path (parity unpinned), so these pin the *oracle*
                 against drift, not against the reference.  Inputs are regenerated from seeds;
                 `digest` guards against RNG drift.
dag_golden.npz   500 synthetic plan DAGs + hand cases with levels / order / status computed by the
                 live NetworkX (the library the reference calls at control_plane.py:104).
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import cases  # noqa: E402
import oracle  # noqa: E402
from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import dag  # noqa: E402

OUT = Path(__file__).resolve().parent
fold = oracle.telemetry_penalty

for name, c in [("config1", cases.config1()), ("small_l2", cases.small_l2(fold)),
                ("edge_cases", cases.edge_cases(fold)), ("config2_small", cases.config2_small(fold))]:
    ids, s64, _ = oracle.topk(c["rows"], c["queries"], c["k"], c["metric"], c["penalty"])
    np.savez_compressed(OUT / f"retrieval_{name}.npz", ids=ids, scores64=s64,
                        digest=cases.digest(c["rows"], c["queries"]))
    print(name, ids.shape)

HAND = [
    # SURVEY.md section 8(c) verified example: generations [[z,y],[w,v,x]]
    {"nodes": [{"name": n} for n in "zyxwv"],
     "edges": [{"from": a, "to": b} for a, b in [("y", "v"), ("z", "w"), ("z", "x"), ("y", "x"), ("z", "v")]]},
    {"nodes": [], "edges": []},                                                        # empty graph
    {"nodes": [{"name": "a"}], "edges": [{"from": "a", "to": "a"}]},                   # self loop
    {"nodes": [{"name": "a"}, {"name": "b"}], "edges": [{"from": "a", "to": "b"}] * 3},  # duplicate edge
    {"nodes": [{"name": "a"}], "edges": [{"from": "q", "to": "a"}, {"from": "a", "to": "r"}]},  # implicit nodes
    {"nodes": [{"name": "a"}, {"name": "b"}, {"name": "a"}, {"name": "c"}],            # duplicate node name
     "edges": [{"from": "c", "to": "a"}, {"from": "a", "to": "b"}]},
    {"nodes": [{"name": n} for n in "abcd"],                                          # cycle behind a prefix
     "edges": [{"from": "a", "to": "b"}, {"from": "b", "to": "c"}, {"from": "c", "to": "b"}, {"from": "c", "to": "d"}]},
]
enc = dag.synthetic_dags(500, seed=4004)
graphs = [dag.to_graph_dict(enc, i) for i in range(len(enc))] + HAND
enc_all = dag.encode_graphs(graphs)
B = len(enc_all)
level = np.full((B, 64), 0xFF, np.uint8)
order = np.full((B, 64), 0xFF, np.uint8)
nlev = np.zeros(B, np.uint8)
status = np.zeros(B, np.uint8)
for i, g in enumerate(graphs):
    lv, od, st = oracle.nx_levels(g)
    idx = {n: v for v, n in enumerate(enc_all.names[i])}
    for n, l in lv.items():
        level[i, idx[n]] = l
    for j, n in enumerate(od):
        order[i, j] = idx[n]
    nlev[i] = (max(lv.values()) + 1) if lv else 0
    status[i] = st
np.savez_compressed(OUT / "dag_golden.npz", n_nodes=enc_all.n_nodes, edge_off=enc_all.edge_off, edges=enc_all.edges,
                    level=level, order=order, n_levels=nlev, status=status)
print("dag", B, int(status.sum()), "cyclic")

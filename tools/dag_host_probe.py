#!/usr/bin/env python
"""K6 through the host-buffer entry point (mcp_dag_levels, pageable numpy arrays in and out) at 100 K and 1 M DAGs."""
import sys, time
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import numpy as np
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as mcp
for n in (100_000, 1_000_000):
    enc = mcp.dag.synthetic_dags(n, seed=4004)
    mcp.dag.levels(enc)
    ts=[]
    for _ in range(5):
        t0=time.perf_counter(); r=mcp.dag.levels(enc); ts.append(time.perf_counter()-t0)
    byts = enc.n_nodes.size*(1+4+64+64+2)+enc.edges.size
    print(n, "host-buffer mcp_dag_levels with order: best %.2f ms median %.2f ms; bytes moved %.1f MB" % (min(ts)*1e3, sorted(ts)[2]*1e3, byts/1e6))

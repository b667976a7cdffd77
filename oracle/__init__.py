"""CPU oracle for the retrieval + level-scheduling hot path.

TEST INFRASTRUCTURE ONLY: imported by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- never by the product package.

Retrieval parity is **unpinned** (the reference holds no ranked query, no tests and no golden
vectors for it, see ``oracle.c``); DAG levels are pinned against NetworkX.
"""
from .cpu_oracle import (  # noqa: F401
    COSINE,
    L2,
    agrees_with_fp32_reference,
    build,
    dag_levels,
    merge,
    nx_levels,
    numpy_topk_baseline,
    numpy_topk_blocked,
    pgvector_f32_scores,
    scores,
    telemetry_penalty,
    topk,
    topk_python,
)

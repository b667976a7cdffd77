"""The retrieval oracle: what it means, and that it has not drifted.

Retrieval parity is *unpinned*: the reference has no ranked query, tests or vectors for it
(SURVEY.md section 8c).  These tests pin the C restatement against (a) a pure-Python statement of the
same arithmetic, (b) hand-derived values for tiny cases, (c) committed golden outputs.
"""
from pathlib import Path

import numpy as np
import pytest

import cases

GOLD = Path(__file__).parent / "golden"


def test_c_matches_pure_python(oracle):
    rng = np.random.default_rng(0)
    e = rng.standard_normal((60, 24)).astype(np.float32)
    e[7] = 0
    e[9] = e[3]
    q = rng.standard_normal((3, 24)).astype(np.float32)
    pen = (rng.random(60) * 0.1).astype(np.float32)
    for metric in (oracle.COSINE, oracle.L2):
        for p in (None, pen):
            ids, s64, s32 = oracle.topk(e, q, 10, metric, p)
            pids, ps = oracle.topk_python(e, q, 10, metric, p)
            assert np.array_equal(ids, pids)
            assert np.array_equal(s64, ps, equal_nan=True)          # bit-for-bit, same summation order
            assert np.array_equal(s32, s64.astype(np.float32), equal_nan=True)


def test_hand_checked_d4(oracle):
    """d = 4, N = 8: orthogonal, identical, opposite, scaled, zero vector, exact ties."""
    q = np.array([[1, 0, 0, 0]], np.float32)
    e = np.array([
        [0, 1, 0, 0],      # 0 orthogonal         cos 1      l2 sqrt2
        [1, 0, 0, 0],      # 1 identical          cos 0      l2 0
        [-1, 0, 0, 0],     # 2 opposite           cos 2      l2 2
        [3, 0, 0, 0],      # 3 parallel, longer   cos 0      l2 2
        [0, 0, 0, 0],      # 4 zero               cos NaN    l2 1
        [1, 0, 0, 0],      # 5 duplicate of 1     cos 0      l2 0
        [1, 1, 0, 0],      # 6 45 degrees         cos 1-1/sqrt2  l2 1
        [0, 0, 2, 0],      # 7 orthogonal         cos 1      l2 sqrt5
    ], np.float32)
    ids, s, _ = oracle.topk(e, q, 8, oracle.COSINE)
    assert ids[0].tolist() == [1, 3, 5, 6, 0, 7, 2, 4]             # ties by id; NaN last
    assert s[0, :3].tolist() == [0.0, 0.0, 0.0]
    assert s[0, 3] == 1.0 - 1.0 / np.sqrt(2.0)
    assert s[0, 4] == 1.0 and s[0, 5] == 1.0 and s[0, 6] == 2.0 and np.isnan(s[0, 7])
    ids, s, _ = oracle.topk(e, q, 8, oracle.L2)
    assert ids[0].tolist() == [1, 5, 4, 6, 0, 2, 3, 7]
    assert s[0].tolist() == [0.0, 0.0, 1.0, 1.0, np.sqrt(2.0), 2.0, 2.0, np.sqrt(5.0)]
    # a penalty reorders: make the duplicate (5) cheaper than the original (1), and 6 worst
    pen = np.array([0, 0.5, 0, 0, 0, 0.25, 9, 0], np.float32)
    ids, s, _ = oracle.topk(e, q, 3, oracle.L2, pen)
    assert ids[0].tolist() == [5, 1, 4] and s[0].tolist() == [0.25, 0.5, 1.0]


def test_k_larger_than_table_and_id_base(oracle):
    e = np.eye(3, 5, dtype=np.float32)
    q = np.ones((1, 5), np.float32)
    ids, s, _ = oracle.topk(e, q, 6, oracle.L2, id_base=1000)
    assert ids[0].tolist() == [1000, 1001, 1002, -1, -1, -1]
    assert np.isinf(s[0, 3:]).all()


@pytest.mark.parametrize("name,make", [
    ("config1", lambda o: cases.config1()),
    ("small_l2", lambda o: cases.small_l2(o.telemetry_penalty)),
    ("edge_cases", lambda o: cases.edge_cases(o.telemetry_penalty)),
    ("config2_small", lambda o: cases.config2_small(o.telemetry_penalty)),
])
def test_golden(oracle, name, make):
    c = make(oracle)
    g = np.load(GOLD / f"retrieval_{name}.npz")
    assert str(g["digest"]) == cases.digest(c["rows"], c["queries"]), "seeded inputs drifted (numpy RNG?)"
    ids, s64, _ = oracle.topk(c["rows"], c["queries"], c["k"], c["metric"], c["penalty"])
    assert np.array_equal(ids, g["ids"])
    assert np.array_equal(s64, g["scores64"], equal_nan=True)


def test_edge_case_semantics(oracle):
    c = cases.edge_cases(oracle.telemetry_penalty)
    ids, s64, _ = oracle.topk(c["rows"], c["queries"], c["k"], c["metric"], None)
    assert ids[0, 0] == 100 and s64[0, 0] < 1e-12                 # identical to the query
    assert ids[0, -1] == 17 and np.isnan(s64[0, -1])               # zero row sorts last
    assert ids[0, -2] == 101 and abs(s64[0, -2] - 2.0) < 1e-12     # opposite vector
    for row in ids:                                                # duplicates 3,40,41,250 tie -> id order
        pos = [int(np.where(row == r)[0][0]) for r in (3, 40, 41, 250)]
        assert pos == list(range(pos[0], pos[0] + 4))
    assert ids[1, 0] == 102                                        # parallel, longer: cosine ignores the norm


def test_merge_equals_global(oracle):
    c = cases.config2_small(oracle.telemetry_penalty, n=1000, q=6, k=20)
    gi, gs, _ = oracle.topk(c["rows"], c["queries"], c["k"], 0, c["penalty"])
    for shards in (2, 3, 8):
        per = -(-1000 // shards)
        li, ls = [], []
        for s in range(shards):
            lo, hi = s * per, min((s + 1) * per, 1000)
            i, sc, _ = oracle.topk(c["rows"][lo:hi], c["queries"], c["k"], 0, c["penalty"][lo:hi], id_base=lo)
            li.append(i)
            ls.append(sc)
        mi, ms = oracle.merge(np.stack(ls), np.stack(li))
        assert np.array_equal(mi, gi) and np.array_equal(ms, gs)


def test_numpy_baseline_agrees_on_separated_data(oracle):
    """The timed BLAS baseline is fp32; on well separated scores it must pick the same ids."""
    c = cases.config2_small(oracle.telemetry_penalty, n=2000, q=4, k=5)
    inv = (1.0 / np.linalg.norm(c["rows"], axis=1)).astype(np.float32)
    ids, _ = oracle.numpy_topk_baseline(c["rows"], inv, c["queries"], 5, 0, c["penalty"])
    gi, _, _ = oracle.topk(c["rows"], c["queries"], 5, 0, c["penalty"])
    assert (ids == gi).mean() > 0.9


@pytest.mark.parametrize("metric", [0, 1])
def test_oracle_within_tolerance_of_pgvector_style_fp32(oracle, metric):
    """north_star's bar: scores within 1e-5 relative of an exact fp32 brute-force reference of pgvector's `<=>` / `<->`.  The
    fp64 oracle is held against pgvector's own arithmetic (float accumulators, oracle.c): every score within 1e-5, and the
    two top-k answers equal except between rows whose fp32 scores are that close."""
    rng = np.random.default_rng(600 + metric)
    rows = rng.standard_normal((4000, 1536), dtype=np.float32)
    rows /= np.linalg.norm(rows, axis=1, keepdims=True)
    rows[100] = rows[7]                                                   # an exact tie
    qs = rng.standard_normal((6, 1536), dtype=np.float32)
    pen = (rng.random(4000) * 0.02).astype(np.float32)
    ids, s64, _ = oracle.topk(rows, qs, 32, metric, pen)
    for i in range(qs.shape[0]):
        ref = oracle.pgvector_f32_scores(rows, qs[i], metric, pen)
        assert np.allclose(oracle.scores(rows, qs[i], metric, pen), ref, rtol=1e-5, atol=0)
        assert oracle.agrees_with_fp32_reference(ids[i], s64[i], ref, 32)
    # and the checker itself notices a wrong answer
    bad = ids[0].copy()
    bad[3] = (bad[3] + 1777) % 4000
    ref0 = oracle.pgvector_f32_scores(rows, qs[0], metric, pen)
    assert not oracle.agrees_with_fp32_reference(bad, s64[0], ref0, 32)

"""One handle over several GPUs (mcp_index_create_multi): the single-process face of the row-sharded registry, the one
``GraphPlanner.plan`` (control_plane.py:57-75) can call.  On a box with one GPU the shards all sit on device 0 -- the fan-out,
the peer copies into the gather buffer and the merge are the same code either way."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _devices():
    import torch

    n = torch.cuda.device_count()
    return list(range(min(n, 8))) if n >= 2 else [0, 0, 0]


@pytest.mark.parametrize("metric", [0, 1])
def test_multi_handle_equals_single_index_and_oracle(mcp, oracle, metric):
    rng = np.random.default_rng(300 + metric)
    n, d, k = 70_003, 128, 24
    rows = rng.standard_normal((n, d), dtype=np.float32)
    rows[n - 2] = rows[7]                                              # a tie across the first and the last shard
    pen = (rng.random(n) * 0.03).astype(np.float32)
    pen[n - 2] = pen[7]
    name = "cosine" if metric == 0 else "l2"
    for q in (1, 5, 200):                                              # streaming path, and the tensor path
        qs = rng.standard_normal((q, d), dtype=np.float32)
        qs[0] = rows[7]
        want_ids, want_s64, _ = oracle.topk(rows, qs, k, metric, pen)
        with mcp.ServiceIndex(d, name, _devices()) as multi, mcp.ServiceIndex(d, name, 0) as one:
            multi.load(rows).set_penalty(pen)
            one.load(rows).set_penalty(pen)
            assert multi.n_rows == n
            mi, ms32, ms64 = multi.search(qs, k)
            oi, _, os64 = one.search(qs, k)
            assert np.array_equal(mi, want_ids) and np.array_equal(ms64, want_s64)
            assert np.array_equal(mi, oi) and np.array_equal(ms64, os64)
            assert np.array_equal(ms32, want_s64.astype(np.float32))
            assert mi[0, 0] == 7 and mi[0, 1] == n - 2
            st = multi.stats()
            assert st["queries"] == q and st["rows_scanned"] == n


def test_multi_handle_updates_ragged_tables_and_refusals(mcp, oracle):
    import torch

    rng = np.random.default_rng(310)
    d = 64
    devs = _devices()
    with mcp.ServiceIndex(d, "cosine", devs) as multi:
        tiny = rng.standard_normal((2, d), dtype=np.float32)           # fewer rows than shards: some shards stay empty
        qs = rng.standard_normal((3, d), dtype=np.float32)
        multi.load(tiny)
        want_ids, want_s64, _ = oracle.topk(tiny, qs, 4, 0, None)
        got = multi.search(qs, 4)
        assert np.array_equal(got[0], want_ids) and np.array_equal(got[2], want_s64, equal_nan=True)
        rows = rng.standard_normal((9_001, d), dtype=np.float32)
        multi.load(rows, id_base=500)
        new = rows.copy()
        new[2990:3015] = qs[0] * np.linspace(1.0, 2.0, 25, dtype=np.float32)[:, None]   # a range that crosses a shard border
        multi.update_rows(2990, new[2990:3015])
        want_ids, want_s64, _ = oracle.topk(new, qs, 8, 0, None, id_base=500)
        got = multi.search(qs, 8)
        assert np.array_equal(got[0], want_ids) and np.array_equal(got[2], want_s64)
        ids = np.array([8_999, 3, 4_500, 2_999, 3_000], np.int64)          # scattered over every shard
        new[ids] = rng.standard_normal((5, d), dtype=np.float32) + qs[1] * 3.0
        multi.upsert_rows(ids, new[ids])
        want_ids, want_s64, _ = oracle.topk(new, qs, 8, 0, None, id_base=500)
        got = multi.search(qs, 8)
        assert np.array_equal(got[0], want_ids) and np.array_equal(got[2], want_s64)
        with pytest.raises(mcp.McpError):
            multi.adopt(torch.zeros((4, d), device="cuda:0"))          # device pointers belong to one GPU
        with pytest.raises(mcp.McpError):
            multi.search_device(torch.zeros((1, d), device="cuda:0"), 2)
        with pytest.raises(mcp.McpError):
            multi.update_rows(9_000, rows[:5])                         # past the end of the table


def test_planner_over_a_sharded_registry(mcp, oracle):
    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import control_plane as cp

    rows = cases.unit_rows(30_000, 256, 91)
    names = [f"s{i}" for i in range(len(rows))]

    class Reg:
        def list_services(self):
            return []

    embed = lambda t: cp.hash_embedding(t, 256)                        # noqa: E731
    planner = cp.GraphPlanner(Reg(), fetch_embeddings=lambda: list(zip(names, rows)), embed=embed, top_k=6, device=_devices())
    planner.refresh_index([])
    intents = [f"intent {i}" for i in range(40)]
    got = planner.retrieve_batch(intents)
    qs = np.stack([embed(t) for t in intents])
    want_ids, _, _ = oracle.topk(rows, qs, 6, 0, np.zeros(len(rows), np.float32))
    assert [[n for n, _ in r] for r in got] == [[names[i] for i in row] for row in want_ids]
    assert planner.retrieve(intents[3]) == got[3]

"""The allocation-free exact search (``mcp_index_search_exact_noalloc_device``) against the ordinary one.

Collected last on purpose (``zz``): a run that stops at the first failure has then already run everything else.
It is for request paths that cannot call cudaMalloc -- a planner thread that would stall another stream's scan,
or a process whose HBM is fully spoken for -- so the test hands in every buffer and only checks answers:
ids must be identical and fp64 scores bit-equal (the tolerance every parity test here uses), on two calls in a
row that share the library's fixed scratch array, plus a repeat of the first call to show that what the second
call left in the array is never read back.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_matches_ordinary_search_twice_in_a_row(mcp):
    import torch

    dev = torch.device("cuda:0")
    rng = np.random.default_rng(4242)
    n, d, k = 30_000, 384, 24
    rows = torch.from_numpy(rng.standard_normal((n, d), dtype=np.float32)).to(dev)
    rows[123] = 0.0                                                    # a NaN score must rank last in both
    rows[5000] = rows[77]                                              # an exact tie must resolve by id in both
    pen = torch.from_numpy((rng.random(n) * 0.03).astype(np.float32)).to(dev)
    pen[5000] = pen[77]                                                # same vector, same penalty: a true tie
    q_first = torch.from_numpy(rng.standard_normal((9, d), dtype=np.float32)).to(dev)
    q_second = torch.from_numpy(rng.standard_normal((64, d), dtype=np.float32)).to(dev)
    q_second[3] = rows[77]
    with mcp.ServiceIndex(d, "cosine", 0) as idx:
        idx.adopt(rows).set_penalty(pen)
        want = {}
        for name, q in (("first", q_first), ("second", q_second)):
            idx.set_option("path", "exact_fp64")
            e = [t.clone() for t in idx.search_device(q, k)]            # existing entry point, fp64 scan
            idx.set_option("path", "auto")
            a = [t.clone() for t in idx.search_device(q, k)]            # existing entry point, default path
            assert idx.sync_status()[1] == 0
            assert torch.equal(e[0], a[0]) and torch.equal(e[2], a[2])
            want[name] = e

        def outputs(nq):                                                # made up front: the call itself allocates nothing
            return (torch.full((nq, k), -7, dtype=torch.int32, device=dev),
                    torch.full((nq, k), -7.0, dtype=torch.float32, device=dev),
                    torch.full((nq, k), -7.0, dtype=torch.float64, device=dev))

        out1, out2, out3 = outputs(9), outputs(64), outputs(9)
        idx.search_exact_noalloc_device(q_first, k, out1)               # first call: also the one-time set-up
        idx.search_exact_noalloc_device(q_second, k, out2)              # second call reuses the same scratch array
        idx.search_exact_noalloc_device(q_first, k, out3)               # stale contents from call 2 must not matter
        torch.cuda.synchronize()
        for got, name in ((out1, "first"), (out2, "second"), (out3, "first")):
            w = want[name]
            assert torch.equal(got[0], w[0]), f"{name}: ids differ"
            assert torch.equal(got[2].isnan(), w[2].isnan())
            assert torch.equal(got[2].nan_to_num(0.0), w[2].nan_to_num(0.0)), f"{name}: fp64 scores differ"
            assert torch.equal(got[1].nan_to_num(0.0), w[1].nan_to_num(0.0))
        assert out2[0][3, 0].item() == 77 and out2[0][3, 1].item() == 5000      # the tie, lowest id first

        # sizes beyond the fixed array are refused with a clear message, not truncated
        with pytest.raises(mcp.McpError) as e:
            idx.search_exact_noalloc_device(torch.zeros((65, d), device=dev), k, outputs(65))
        assert e.value.code == -1 and "64" in str(e.value)


def test_noalloc_refuses_tables_beyond_the_fixed_array(mcp):
    import torch

    dev = torch.device("cuda:0")
    n, d = (1 << 20) + 1, 8
    rows = torch.ones((n, d), dtype=torch.float32, device=dev)
    with mcp.ServiceIndex(d, "l2", 0) as idx:
        idx.adopt(rows)
        out = (torch.empty((1, 4), dtype=torch.int32, device=dev), torch.empty((1, 4), dtype=torch.float32, device=dev),
               torch.empty((1, 4), dtype=torch.float64, device=dev))
        with pytest.raises(mcp.McpError) as e:
            idx.search_exact_noalloc_device(torch.ones((1, d), device=dev), 4, out)
        assert e.value.code == -1 and "1048576" in str(e.value)

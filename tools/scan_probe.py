import sys, torch
sys.path.insert(0, "/root/repo")
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg
import bench
dev = torch.device("cuda:0")
n, dim, k = 200_000, 1536, 32
rr = bench.make_shard(torch, dev, 0, n, dim, 5)
for nq in (1, 8):
    qr = bench.make_queries(torch, dev, nq, dim, 6)
    idx = pkg.ServiceIndex(dim, "cosine", 0).adopt(rr)
    idx.set_option("path", 3)
    for _ in range(3):
        idx.search_device(qr, k)
    torch.cuda.synchronize()
    ms = bench.time_steps(torch, lambda: idx.search_device(qr, k), 10)
    print(nq, "queries:", ms * 1e3, "us")
    idx.close()

import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import time, numpy as np, torch
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as mcp
dev=torch.device('cuda:0')
n,d,q,k=100000,1536,256,32
rows=torch.randn((n,d),device=dev); qs=torch.randn((q,d),device=dev)
idx=mcp.ServiceIndex(d,"cosine"); idx.adopt_device(rows) if hasattr(idx,'adopt_device') else idx.adopt(rows)
for _ in range(5): out=idx.search_device(qs,k)
torch.cuda.synchronize()
t0=time.perf_counter()
for _ in range(50): out=idx.search_device(qs,k)
t1=time.perf_counter()
torch.cuda.synchronize()
t2=time.perf_counter()
print("host enqueue per call us", (t1-t0)/50*1e6, " total per call us", (t2-t0)/50*1e6)
qh=qs.cpu().numpy()
for _ in range(3): idx.search(qh,k)
t0=time.perf_counter()
for _ in range(50): idx.search(qh,k)
t1=time.perf_counter()
print("host-buffer search per call us", (t1-t0)/50*1e6)

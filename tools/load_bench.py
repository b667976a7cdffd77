#!/usr/bin/env python
"""Table-file load speed: writes an MCPTBL01 file of --rows x 1536 fp32, reads it once with dd
(what the file system delivers: page cache or disk, whichever it is on this box), then loads it through
mcp_index_load_file (pinned chunk ring, two reader threads, one cudaMemcpyAsync per chunk) and through the old route
(np.memmap -> pageable cudaMemcpy2DAsync).  Prints GB/s for each.

  python tools/load_bench.py [--rows 1000000] [--dir /tmp] [--drop-caches]
"""
import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000)
ap.add_argument("--dim", type=int, default=1536)
ap.add_argument("--dir", default="/tmp")
a = ap.parse_args()
path = Path(a.dir) / "mcp_load_bench.tbl"
rng = np.random.default_rng(1)
hdr = np.zeros(8, np.int64)
hdr[0] = int.from_bytes(b"MCPTBL01", "little")
hdr[1:5] = (a.dim, a.rows, 0, 0)
with open(path, "wb") as f:
    f.write(hdr.tobytes())
    block = rng.standard_normal((50_000, a.dim), dtype=np.float32)
    for lo in range(0, a.rows, 50_000):
        f.write(block[: min(50_000, a.rows - lo)].tobytes())
gb = (a.rows * a.dim * 4) / 1e9
out = {"file_GB": round(gb, 2)}
for rep in range(2):
    t0 = time.perf_counter()
    subprocess.run(["dd", f"if={path}", "of=/dev/null", "bs=32M"], check=True, capture_output=True)
    out[f"dd_GBps_pass{rep + 1}"] = round(gb / (time.perf_counter() - t0), 2)
with pkg.ServiceIndex(a.dim, "cosine", 0) as idx:
    idx.set_option("shadow", 0)
    for rep in range(2):
        t0 = time.perf_counter()
        idx.load_file(path)
        out[f"load_file_GBps_pass{rep + 1}"] = round(gb / (time.perf_counter() - t0), 2)
    t0 = time.perf_counter()
    rows = np.memmap(path, dtype=np.float32, mode="r", offset=64, shape=(a.rows, a.dim))
    idx.load(rows)
    out["memmap_load_GBps"] = round(gb / (time.perf_counter() - t0), 2)
with pkg.ServiceIndex(a.dim, "cosine", 0) as idx:
    t0 = time.perf_counter()
    idx.load_file(path)
    out["load_file_with_shadow_GBps"] = round(gb / (time.perf_counter() - t0), 2)
os.unlink(path)
print(json.dumps(out))

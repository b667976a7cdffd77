"""In-tree build of the CUDA library: one nvcc call, sm_100a only.

``python -m autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200.build``
or ``__graft_entry__.build()``.  nvcc cross-compiles without a GPU; the resulting
``libmcp_b200.so`` sits beside this file (git-ignored, but it travels with the tree).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
SRC = PKG / "csrc" / "mcp_abi.cu"
OUT = PKG / "libmcp_b200.so"
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--compiler-options", "-fPIC",
    "-shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: the CUDA library cannot be built (there is no CPU fallback)")


def sources():
    return sorted((PKG / "csrc").glob("*.cu*")) + [PKG.parent / "include" / "mcp_b200.h"]


def is_stale() -> bool:
    if not OUT.exists():
        return True
    t = OUT.stat().st_mtime
    return any(p.stat().st_mtime > t for p in sources())


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not is_stale():
        return OUT
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", str(OUT), str(SRC)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed ({res.returncode}):\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr, file=sys.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))

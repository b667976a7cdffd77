#!/usr/bin/env python
"""bench.py -- top-k retrieval queries/sec on B200 (BASELINE.json metric), one process per GPU.

  python bench.py --gpus 1 --steps K --warmup W            config 2: 100K x 1536, 256 queries, top-32 cosine + telemetry
  torchrun ... bench.py --gpus N --steps K --warmup W      config 3: 10M x 1536 row-sharded over N GPUs, 1024 queries,
                                                           top-64, NCCL all-gather merge (strong scaling: the table is fixed)
  python bench.py --impl reference ...                     the CPU path (numpy restatement of the pgvector exact scan,
                                                           BASELINE.md section 3) on the host cores, same config

A "step" is one batch of queries answered over the whole table through the C ABI.  `value` is timed with
CUDA events with everything resident in HBM; `e2e` is the same metric through the host-buffer API with the
H2D / D2H copies inside the timed region.  One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

CONFIGS = {
    # BASELINE.json configs[1]
    "config2": dict(n=100_000, q=256, k=32, dim=1536, metric="cosine", seeds=(1002, 2002, 3002),
                    label="100K services x 1536-d fp32, batch 256, top-32 cosine + telemetry penalty"),
    # BASELINE.json configs[2]
    "config3": dict(n=10_000_000, q=1024, k=64, dim=1536, metric="cosine", seeds=(1003, 2003, 3003),
                    label="10M services x 1536-d fp32 (61.4 GB) row-sharded, batch 1024, top-64 cosine + telemetry penalty"),
}
METRIC_NAME = "top-k retrieval queries/sec"
UNIT = "queries/s"


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sus=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    sm_max=float(d.get("sm_max_mhz", 1965.0)), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, sm_max=1965.0, source="fallback (B200_PROFILING.md)")


# ---------------------------------------------------------------------------------------------
# clocks during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi polled every 20 ms from before the warm-up on (it needs a few hundred ms to start, longer than a
    short timed region); summary() keeps the samples that fall inside the window it is given."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.monotonic(), [c.strip() for c in line.split(",")]))

    def count(self, t0, t1):
        return sum(1 for t, r in self.rows if t0 <= t <= t1 and len(r) >= 8)

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(2)
            except Exception:  # noqa: BLE001
                self.proc.kill()

    def summary(self, t0, t1, window):
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 8]
        sm = sorted(float(r[1]) for r in rows if r[1].replace(".", "").isdigit())
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "window": window}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[4 + i].lower().startswith("active") for r in rows)]
        mx = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm),
                "window": window}


# ---------------------------------------------------------------------------------------------
# synthetic data on the device (no dataset offline): unit-norm standard-normal rows, SURVEY.md section 8d telemetry
# ---------------------------------------------------------------------------------------------
def measure_tf32_peak(torch, dev):
    """cuBLAS TF32 8192^3 on this GPU, this run: (burst TFLOP/s = best of 5, sustained = 1 s loop).
    MEASURED_PEAKS.json only records bf16; north_star asks for the tensor-pipe fraction of a TF32 kernel."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn((n, n), device=dev)
        b = torch.randn((n, n), device=dev)
        for _ in range(2):
            a @ b
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(4, int(1000.0 / best))
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record()
        torch.cuda.synchronize()
        sus = e0.elapsed_time(e1) / reps
        fl = 2.0 * n ** 3 / 1e12
        return fl / (best * 1e-3), fl / (sus * 1e-3)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def make_shard(torch, dev, lo, hi, dim, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    rows = torch.empty((hi - lo, dim), dtype=torch.float32, device=dev)
    step = 262_144
    for a in range(0, hi - lo, step):
        b = min(a + step, hi - lo)
        torch.randn((b - a, dim), generator=g, device=dev, out=rows[a:b])
        rows[a:b] /= rows[a:b].norm(dim=1, keepdim=True)
    return rows


def make_penalty(torch, dev, n, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    lat = torch.exp(torch.randn(n, generator=g, device=dev) * 0.5 + 3.912023).clamp_(1.0, 2000.0)   # LogNormal(ln 50, 0.5)
    u = torch.rand(n, generator=g, device=dev)
    err = 1.0 - (1.0 - u) ** (1.0 / 50.0)                                                          # Beta(1, 50) by inversion
    cost = torch.rand(n, generator=g, device=dev) * 0.019 + 0.001
    return (0.05 * torch.clamp(lat / 1000.0, max=1.0) + 0.5 * err + 1.0 * cost).float().contiguous()


def make_queries(torch, dev, q, dim, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    x = torch.randn((q, dim), generator=g, device=dev)
    return (x / x.norm(dim=1, keepdim=True)).contiguous()


# ---------------------------------------------------------------------------------------------
# the CPU arm (numpy restatement; pgvector / Postgres do not exist on either machine)
# ---------------------------------------------------------------------------------------------
def cpu_arm(cfg, steps, warmup, sample_rows=None, rows=None, queries=None, penalty=None):
    """-> dict(value=q/s, ms_per_step, cores, sample).  Times `steps` batches on a bounded row sample and scales
    linearly in rows to the config's table (the scan is linear in N)."""
    import numpy as np

    import oracle

    n, q, k, dim = cfg["n"], cfg["q"], cfg["k"], cfg["dim"]
    m = min(n, sample_rows or n)
    if rows is None:
        rng = np.random.default_rng(cfg["seeds"][0])
        rows = rng.standard_normal((m, dim), dtype=np.float32)
        rows /= np.linalg.norm(rows, axis=1, keepdims=True)
        queries = np.random.default_rng(cfg["seeds"][1]).standard_normal((q, dim), dtype=np.float32)
        penalty = oracle.telemetry_penalty(*[np.abs(np.random.default_rng(cfg["seeds"][2] + i).standard_normal(m)).astype(np.float32) * s
                                             for i, s in enumerate((50.0, 0.02, 0.01))])
    rows = rows[:m]
    penalty = penalty[:m]
    inv = (1.0 / np.linalg.norm(rows, axis=1)).astype(np.float32)
    mid = 0 if cfg["metric"] == "cosine" else 1
    n2 = np.einsum("ij,ij->i", rows, rows).astype(np.float32) if mid else None
    for _ in range(warmup):
        oracle.numpy_topk_baseline(rows, inv, queries, k, mid, penalty, n2)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.numpy_topk_baseline(rows, inv, queries, k, mid, penalty, n2)
    dt = (time.perf_counter() - t0) / steps
    scale = n / m
    return dict(value=q / (dt * scale), ms_per_step=dt * scale * 1e3, cores=oracle.cpu_oracle.cpu_count(),
                sample=(f"{steps} batches of {q} queries over {m} of {n} rows" + (f", time scaled x{scale:g} (scan is linear in rows)" if m < n else "")
                        + "; numpy/OpenBLAS sgemm + argpartition, all host threads"))


def run_reference(args, world, rank):
    if rank != 0:
        return
    cfg = CONFIGS["config2" if (args.workload == "auto" and world == 1) or args.workload == "config2" else "config3"]
    cfg = override(cfg, args)
    sample = 100_000 if cfg["n"] > 100_000 else None
    r = cpu_arm(cfg, max(1, min(args.steps, 5)), min(args.warmup, 1), sample_rows=sample)
    line = {
        "impl": "reference", "metric": METRIC_NAME, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["label"], "rows": cfg["n"], "queries": cfg["q"], "k": cfg["k"], "dim": cfg["dim"]},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference has no ranked query and no pgvector here (SURVEY.md section 0); this is BASELINE.md section 3's numpy restatement",
    }
    print(json.dumps(line), flush=True)


def override(cfg, args):
    cfg = dict(cfg)
    for key in ("n", "q", "k"):
        v = getattr(args, {"n": "rows", "q": "queries", "k": "topk"}[key])
        if v:
            cfg[key] = v
            cfg["label"] += f" [{key}={v} override]"
    if getattr(args, "metric", None):
        cfg["metric"] = args.metric
        cfg["label"] = cfg["label"].replace("cosine", args.metric) + f" [metric={args.metric} override]"
    return cfg


# ---------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "config2", "config3"])
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--queries", type=int, default=0)
    ap.add_argument("--topk", type=int, default=0)
    ap.add_argument("--metric", choices=["cosine", "l2"], default=None, help="override the workload's distance (SURVEY 8d: config 3 is also run with L2)")
    ap.add_argument("--path", default="auto", choices=["auto", "stream_fp32", "gemm_tf32", "gemm_f16"])
    ap.add_argument("--opt", action="append", default=[], help="library option name=value (repeatable), e.g. gemm_wide=0")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--dag", type=int, default=100_000, help="also time the Kahn-level kernel on this many synthetic DAGs (extra; 0 = skip)")
    ap.add_argument("--no-config3-ref", action="store_true", help="skip the single-GPU run of the multi-GPU workload (extra)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, world, rank)

    import numpy as np
    import torch
    import torch.distributed as dist

    import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg
    from autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 import sharded

    assert torch.cuda.is_available(), "bench.py needs a GPU: the product path has no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    saved_stdout = None
    if world > 1:
        # NCCL prints its version banner (and everything NCCL_DEBUG asks for) on stdout while the communicator comes up;
        # stdout must carry exactly one JSON line, so fd 1 points at stderr until the first collectives are through
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    wl = "config2" if (args.workload == "auto" and world == 1) or args.workload == "config2" else "config3"
    cfg = override(CONFIGS[wl], args)
    n, q, k, dim = cfg["n"], cfg["q"], cfg["k"], cfg["dim"]
    pk = peaks()

    lo, hi = sharded.shard_bounds(n, world, rank)
    rows = make_shard(torch, dev, lo, hi, dim, cfg["seeds"][0] + 7919 * rank)
    pen = make_penalty(torch, dev, hi - lo, cfg["seeds"][2] + 7919 * rank)
    qs = make_queries(torch, dev, q, dim, cfg["seeds"][1])
    sh = sharded.ShardedIndex(dim, cfg["metric"], local)
    sh.adopt_shard(rows, n).set_penalty(pen)
    if args.path != "auto":
        sh.local.set_option("path", args.path)
    for o in args.opt:
        name, val = o.split("=")
        sh.local.set_option(name, int(val))
    sh.local.set_option("time_coarse", 1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local)
    clocks.__enter__()
    for _ in range(args.warmup):
        out = sh.search_device(qs, k)
    barrier()
    sh.sync_status()
    if saved_stdout is not None:
        t_ = torch.zeros(1, device=dev)
        dist.all_reduce(t_)                      # every kind of collective used later has run once by now
        torch.cuda.synchronize()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    st0 = sh.local.stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    tw0 = time.monotonic()
    e0.record()
    for _ in range(args.steps):
        out = sh.search_device(qs, k)
    e1.record()
    barrier()
    tw1 = time.monotonic()
    ms_total = e0.elapsed_time(e1)
    fails, unresolved = sh.sync_status()
    st1 = sh.local.stats()
    t = torch.tensor([ms_total], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    clock_window = "timed region"
    need = torch.tensor([1.0 if clocks.count(tw0, tw1) < 3 else 0.0], device=dev)
    if world > 1:
        dist.all_reduce(need, op=dist.ReduceOp.MAX)
    if need.item() > 0:
        # the timed region is shorter than a few nvidia-smi periods: keep the same step loop running (untimed, the same
        # number of steps on every rank) until the sampler has seen it under load, and report those samples
        n_probe = max(args.steps, min(20000, int(500.0 / ms_step) + 1))
        tw0 = time.monotonic()
        for _ in range(n_probe):
            sh.search_device(qs, k)
        barrier()
        tw1 = time.monotonic()
        sh.sync_status()
        clock_window = "the same step loop continued untimed for about 0.5 s (timed region shorter than the sampling period)"
    clocks.__exit__(None, None, None)
    clock_summary = clocks.summary(tw0, tw1, clock_window)
    value = q / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel (the coarse scan), timed by CUDA events inside the library on the launch stream
    launches = st1["kernel_launches"] - st0["kernel_launches"] + (args.steps if world > 1 else 0)   # + K5 merge per step
    coarse_n = max(1, st1["coarse_timed"] - st0["coarse_timed"])
    coarse_ms = (st1["coarse_ms"] - st0["coarse_ms"]) / coarse_n
    n_loc = hi - lo
    alg_bytes = n_loc * dim * 4 + n_loc * 8 + q * dim * 4 + q * k * 8
    flops = 2.0 * q * n_loc * dim
    path = st1["last_path"]
    hbm_ach = alg_bytes / (coarse_ms * 1e-3) / 1e9 if coarse_ms > 0 else 0.0
    tf_ach = flops / (coarse_ms * 1e-3) / 1e12 if coarse_ms > 0 else 0.0
    # cuBLAS TF32 is measured at the very end of the run (it heats the part); until then use half of bf16
    tf32_burst, tf32_sus = pk["bf16"] / 2.0, pk["bf16_sus"] / 2.0
    tf32_peak = tf32_sus if coarse_ms > 2.0 else tf32_burst
    traffic = None
    tp = ROOT / "profiles" / "traffic.json"
    if tp.exists():
        traffic = json.loads(tp.read_text()).get(f"{wl}_path{path}")
    hbm_roof = {"bound": "hbm", "achieved": round(hbm_ach, 1), "peak": pk["hbm"], "unit": "GB/s", "frac": round(hbm_ach / pk["hbm"], 4),
                "traffic": traffic, "kernel": {1: "scan_fp32_kernel", 2: "gemm_tf32_kernel", 3: "exact_scan_kernel"}.get(path),
                "kernel_ms": round(coarse_ms, 4), "algorithmic_bytes": alg_bytes, "peak_source": pk["source"]}
    tensor_roof = {"bound": "tensor", "achieved": round(tf_ach, 2), "peak": round(tf32_peak, 1), "unit": "TFLOP/s",
                   "frac": round(tf_ach / tf32_peak, 4), "traffic": traffic, "kernel": hbm_roof["kernel"], "kernel_ms": round(coarse_ms, 4),
                   "flops": flops, "peak_source": "half of measured bf16"}
    # the bound is whichever floor is higher for this launch (SURVEY.md section 7.2a ridge)
    use_tensor = path == 2 and flops / (tf32_peak * 1e12) > alg_bytes / (pk["hbm"] * 1e9)
    roofline, roofline_alt = (tensor_roof, hbm_roof) if use_tensor else (hbm_roof, tensor_roof)

    # ---- e2e: host buffers through the public API, copies inside the timed region
    q_host = torch.empty((q, dim), dtype=torch.float32, pin_memory=True)
    q_host.copy_(qs)
    e2e_steps = max(3, min(args.steps, 10))
    h2d, d2h = q * dim * 4, q * k * 16
    if world == 1:
        qn = q_host.numpy()
        for _ in range(2):
            sh.local.search(qn, k)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ids_h, s_h, s64_h = sh.local.search(qn, k)
        e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
    else:
        outs_h = [torch.empty((q, k), dtype=dt, pin_memory=True) for dt in (torch.int32, torch.float32, torch.float64)]

        def e2e_step():
            qd = q_host.to(dev, non_blocking=True)
            res = sh.search_device(qd, k)
            for o, r in zip(outs_h, res):
                o.copy_(r, non_blocking=True)
            torch.cuda.synchronize()

        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e = {"value": q / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": round(e2e_ms, 4), "api": "ServiceIndex.search (mcp_index_search, host buffers)" if world == 1
           else "pinned host -> ShardedIndex.search_device -> pinned host"}

    extra = {"certificate_failures": fails, "unresolved": unresolved, "kprime": st1["last_kprime"], "path": path,
             "roofline_alt": roofline_alt}
    cpu_baseline = None
    if rank == 0 and world == 1:
        if not args.no_extras:
            sweep = {}
            for b in (1, 8, 64, 256, 1024):
                qb = make_queries(torch, dev, b, dim, 99 + b)
                for _ in range(3):
                    sh.local.search_device(qb, k)
                torch.cuda.synchronize()
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                for _ in range(10):
                    sh.local.search_device(qb, k)
                a1.record()
                torch.cuda.synchronize()
                ms = a0.elapsed_time(a1) / 10
                sweep[str(b)] = {"ms": round(ms, 4), "qps": round(b / ms * 1e3, 1),
                                 "table_GBps": round(n * dim * 4 / (ms * 1e-3) / 1e9, 1)}
            sh.sync_status()
            extra["batch_sweep_whole_search"] = sweep
        if args.dag:
            enc = pkg.dag.synthetic_dags(args.dag, seed=4004)
            nn = torch.from_numpy(enc.n_nodes).to(dev)
            off = torch.from_numpy(enc.edge_off.view(np.int32)).to(dev)
            ed = torch.from_numpy(enc.edges).to(dev).contiguous()
            outb = pkg.dag.levels_device(nn, off, ed)
            torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            for _ in range(10):
                pkg.dag.levels_device(nn, off, ed, out=outb)
            a1.record()
            torch.cuda.synchronize()
            ms = a0.elapsed_time(a1) / 10
            byts = enc.n_nodes.size * (1 + 4 + 128 + 2) + enc.edges.size
            extra["dag_levels"] = {"dags": args.dag, "ms": round(ms, 4), "dags_per_s": round(args.dag / ms * 1e3),
                                   "GBps": round(byts / (ms * 1e-3) / 1e9, 1), "with_order": True}
        if wl == "config2" and not args.no_config3_ref and not args.no_extras and not (args.rows or args.queries or args.topk):
            # the multi-GPU workload (config 3) on this one GPU, so that the N = 2, 4, 8 lines (strong scaling over a fixed
            # 10M-row table) have an N = 1 point; the headline N = 1 line above is config 2 as BASELINE.json quotes it
            try:
                c3 = CONFIGS["config3"]
                sh.close()
                del rows, pen
                torch.cuda.empty_cache()
                rows3 = make_shard(torch, dev, 0, c3["n"], dim, c3["seeds"][0])
                pen3 = make_penalty(torch, dev, c3["n"], c3["seeds"][2])
                q3 = make_queries(torch, dev, c3["q"], dim, c3["seeds"][1])
                sh = sharded.ShardedIndex(dim, c3["metric"], local)
                sh.adopt_shard(rows3, c3["n"]).set_penalty(pen3)
                for _ in range(2):
                    sh.search_device(q3, c3["k"])
                torch.cuda.synchronize()
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                for _ in range(5):
                    sh.search_device(q3, c3["k"])
                a1.record()
                torch.cuda.synchronize()
                ms3 = a0.elapsed_time(a1) / 5
                f3, u3 = sh.sync_status()
                extra["config3_on_1_gpu"] = {"workload": c3["label"], "ms_per_step": round(ms3, 3), "queries_per_s": round(c3["q"] / ms3 * 1e3, 1),
                                             "TFLOPs": round(2.0 * c3["q"] * c3["n"] * dim / (ms3 * 1e-3) / 1e12, 1),
                                             "certificate_failures": f3, "unresolved": u3}
                rows, pen, qs = rows3[:n], pen3[:n].contiguous(), make_queries(torch, dev, q, dim, cfg["seeds"][1])   # for the CPU baseline below
            except Exception as ex:  # noqa: BLE001 -- an extra must never take the headline line down
                extra["config3_on_1_gpu"] = {"error": str(ex)[:200]}
        if not args.no_cpu_baseline:
            import oracle

            rows_np, qs_np, pen_np = rows.cpu().numpy(), qs.cpu().numpy(), pen.cpu().numpy()
            sample = 100_000 if n > 100_000 else None
            cb = cpu_arm(cfg, 3, 1, sample_rows=sample, rows=rows_np, queries=qs_np, penalty=pen_np)
            cpu_baseline = {"value": cb["value"], "unit": UNIT, "cores": cb["cores"], "kind": "port", "sample": cb["sample"]}
            del rows_np

    if rank == 0:
        # the tensor denominator, measured now that every timed region is over
        tf32_burst, tf32_sus = measure_tf32_peak(torch, dev)
        # a kernel timed inside a long, power-capped step is held against the sustained figure, a short one against the burst
        tf32_peak = tf32_sus if coarse_ms > 2.0 else tf32_burst
        for rf in (roofline, roofline_alt):
            if rf["bound"] == "tensor":
                rf["peak"] = round(tf32_peak, 1)
                rf["frac"] = round(tf_ach / tf32_peak, 4)
                rf["peak_source"] = (f"cuBLAS TF32 8192^3 measured at the end of this run: burst {tf32_burst:.0f}, sustained {tf32_sus:.0f} "
                                     "TFLOP/s (MEASURED_PEAKS.json holds bf16 only)")
        line = {
            "metric": METRIC_NAME, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if wl == "config3" else "weak",
            "vs_baseline": None, "dtype": "f32 table; tf32 tensor-core coarse pass + f64 exact rescoring" if path == 2 else "f32 coarse pass + f64 exact rescoring",
            "data": "synthetic",
            "config": {"workload": cfg["label"], "rows": n, "rows_per_gpu": n_loc, "queries": q, "k": k, "dim": dim,
                       "parallelism": f"row-shard x{world}" + (" + NCCL all-gather merge" if world > 1 else ""),
                       "l2": f"table shard {n_loc * dim * 4 / 1e6:.0f} MB > 126 MB L2, streamed once per step (no flush needed)"},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clock_summary, "extra": extra,
        }
        print(json.dumps(line), flush=True)
    sh.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

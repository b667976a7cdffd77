"""Telemetry fold: (latency, error rate, cost_profile) -> one fp32 penalty per service.

The reference names the fields but holds no formula and no Redis key schema
(``cost_profile``: control_plane.py:31 and README.md:93; latency / error rate: README.md:43-48),
so the formula is this build's:

    penalty = w_lat * min(latency_ms / lat_ref, 1) + w_err * error_rate + w_cost * cost_profile

computed once per registry refresh, in fp32, on the host, and handed to
``ServiceIndex.set_penalty``; the kernels only ever see the folded array.  Weights come from
environment variables in the style of control_plane.py:17-19.
"""
from __future__ import annotations

import json
import os
from typing import Iterable, Optional

import numpy as np

TELEMETRY_PREFIX = "mcp:telemetry:"   # beside SERVICES_PREFIX = "mcp:service:" (control_plane.py:20)


def weights_from_env() -> dict:
    return {
        "w_lat": float(os.getenv("MCP_W_LATENCY", "0.05")),
        "w_err": float(os.getenv("MCP_W_ERROR", "0.5")),
        "w_cost": float(os.getenv("MCP_W_COST", "1.0")),
        "lat_ref": float(os.getenv("MCP_LATENCY_REF_MS", "1000")),
    }


def fold_penalty(latency_ms, error_rate, cost_profile, w_lat=0.05, w_err=0.5, w_cost=1.0, lat_ref=1000.0) -> np.ndarray:
    lat = np.asarray(latency_ms, np.float32)
    err = np.asarray(error_rate, np.float32)
    cost = np.asarray(cost_profile, np.float32)
    pen = (np.float32(w_lat) * np.minimum(lat / np.float32(lat_ref), np.float32(1.0))
           + np.float32(w_err) * err + np.float32(w_cost) * cost)
    return pen.astype(np.float32)


def penalty_for_services(services: Iterable[dict], redis_client=None, **weights) -> np.ndarray:
    """Registry records (README.md:86-95 shape) -> penalty array in record order.

    ``cost_profile`` comes from the record; latency / error rate from ``mcp:telemetry:<name>``
    (JSON ``{"latency_ms": .., "error_rate": ..}``) when a Redis client is given, else 0.
    """
    lat, err, cost = [], [], []
    for s in services:
        t = {}
        if redis_client is not None:
            raw = redis_client.get(f"{TELEMETRY_PREFIX}{s['name']}")
            if raw:
                t = json.loads(raw)
        lat.append(float(t.get("latency_ms", 0.0)))
        err.append(float(t.get("error_rate", 0.0)))
        cp = s.get("cost_profile", 0.0)
        cost.append(float(cp if not isinstance(cp, dict) else cp.get("cost", 0.0)))
    w = {**weights_from_env(), **weights}
    return fold_penalty(lat, err, cost, **w)

"""Seeded input generators shared by the golden-vector script and the tests."""
import hashlib

import numpy as np


def unit_rows(n, d, seed):
    x = np.random.default_rng(seed).standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True).astype(np.float32)
    return x


def telemetry(n, seed):
    """config 2/3 telemetry distribution -> (latency_ms, error_rate, cost_profile)."""
    rng = np.random.default_rng(seed)
    lat = np.clip(rng.lognormal(np.log(50.0), 0.5, n), 1.0, 2000.0).astype(np.float32)
    err = rng.beta(1.0, 50.0, n).astype(np.float32)
    cost = rng.uniform(0.001, 0.02, n).astype(np.float32)
    return lat, err, cost


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def config1():
    """1K x 1536, one query, top-8 cosine, no penalty."""
    return dict(rows=unit_rows(1000, 1536, 1001), queries=unit_rows(1, 1536, 2001), k=8, metric=0, penalty=None)


def small_l2(fold):
    rows = np.random.default_rng(11).standard_normal((2000, 96), dtype=np.float32) * 3.0   # un-normalised
    q = np.random.default_rng(12).standard_normal((5, 96), dtype=np.float32)
    return dict(rows=rows, queries=q, k=16, metric=1, penalty=fold(*telemetry(2000, 13)))


def edge_cases(fold):
    """duplicates (exact ties -> id order), a zero row (cosine NaN -> last), un-normalised rows, D = 1530."""
    rng = np.random.default_rng(21)
    rows = rng.standard_normal((300, 1530), dtype=np.float32) * rng.uniform(0.1, 10.0, (300, 1)).astype(np.float32)
    q = rng.standard_normal((4, 1530), dtype=np.float32)
    rows[17] = 0.0
    rows[40] = rows[3]; rows[41] = rows[3]; rows[250] = rows[3]
    rows[100] = q[0]                      # identical to a query: distance exactly 0
    rows[101] = -q[0]                     # opposite: cosine distance 2
    rows[102] = 2.5 * q[1]                # parallel, different norm
    return dict(rows=rows, queries=q, k=300, metric=0, penalty=None)


def config2_small(fold, n=4096, q=32, k=32):
    """config-2 shaped (cosine + telemetry penalty) at a size the oracle does in a second."""
    return dict(rows=unit_rows(n, 1536, 1002), queries=unit_rows(q, 1536, 2002), k=k, metric=0,
                penalty=fold(*telemetry(n, 3002)))

"""Drop-in mirror of the reference control plane for the hot path only.

Same names, request / response models, endpoints and error behaviour as the reference
(``ServiceRegistry``, ``GraphPlanner.plan``, ``Orchestrator.execute``, ``POST /plan``,
``/execute``, ``/plan_and_execute``: control_plane.py:26-151).  What changes is *inside* two seams:

* ``GraphPlanner.plan`` (control_plane.py:57-75) no longer lists every registered service in
  the prompt.  Between the registry read (:58) and the prompt loop (:65-66) it embeds the intent,
  asks the HBM-resident ``ServiceIndex`` for the top-k services by pgvector distance + telemetry
  penalty, and prompts with those k candidates only.
* ``Orchestrator.execute`` (control_plane.py:93-131) takes its node order from the batched Kahn
  kernel instead of ``nx.topological_sort`` (:104); ``schedule_batch`` exposes the batched form
  for planners that score many candidate DAGs.

Everything else the reference does around those seams (Redis scan, OpenAI chat call, HTTP
fan-out) is network I/O and out of scope; it is kept injectable so the tests can
stub it.  The app is built by ``create_app`` instead of at import time: the reference connects
to Postgres while importing (:137), which here would mean allocating the table on the GPU at
import.

Deliberate differences from the reference, each visible in a test:
  * the fallback lookup ``G.edges[in_edges[0]][name]`` (:119) always raises ``KeyError`` in the
    reference, so a failing node with an in-edge answers 500; here the first in-edge's
    ``fallback`` attribute is read as evidently intended.
  * ``/plan_and_execute`` runs the blocking ``plan`` in a worker thread instead of on the event
    loop (:150).
"""
from __future__ import annotations

import asyncio
import hashlib
import json
import logging
import os
import threading
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
from pydantic import BaseModel

from . import dag as _dag
from . import telemetry as _telemetry
from .index import ServiceIndex

REDIS_URL = os.getenv("REDIS_URL", "redis://localhost:6379/0")
POSTGRES_DSN = os.getenv("POSTGRES_DSN", "host=localhost dbname=metadata user=postgres password=secret")
OPENAI_API_KEY = os.getenv("OPENAI_API_KEY")
SERVICES_PREFIX = "mcp:service:"
# knobs this build adds, same style as the three above
RETRIEVAL_METRIC = os.getenv("MCP_RETRIEVAL_METRIC", "cosine")
RETRIEVAL_TOP_K = int(os.getenv("MCP_RETRIEVAL_TOP_K", "8"))
EMBED_DIM = int(os.getenv("MCP_EMBED_DIM", "1536"))
_dev = os.getenv("MCP_CUDA_DEVICE", "0")
CUDA_DEVICE = [int(d) for d in _dev.split(",")] if "," in _dev else int(_dev)   # "0,1,2,3": the registry row-sharded over four GPUs


# ---- registry (control_plane.py:26-35) --------------------------------------------------------
class ServiceRegistry:
    """``list_services`` as in the reference; the Redis client may be injected (tests use a dict-backed fake)."""

    def __init__(self, redis_url: str = REDIS_URL, client=None):
        if client is None:
            import redis  # the reference's dependency; absent in this image

            client = redis.from_url(redis_url)
        self.redis = client

    def list_services(self) -> List[dict]:
        return [json.loads(self.redis.get(key)) for key in self.redis.scan_iter(f"{SERVICES_PREFIX}*")]

    def get_services(self, names: Sequence[str]) -> List[dict]:
        """The k candidates' records only: k GETs instead of SCAN + one GET per registered service (the
        reference's :33-34 is O(N) round trips per request, unworkable at 1M services).  Unknown names are skipped."""
        out = []
        for n in names:
            raw = self.redis.get(f"{SERVICES_PREFIX}{n}")
            if raw:
                out.append(json.loads(raw))
        return out


# ---- wire models (control_plane.py:39-43, :79-85) -----------------------------------------------
class PlanRequest(BaseModel):
    intent: str


class PlanResponse(BaseModel):
    graph: dict


class ExecuteRequest(BaseModel):
    graph: dict
    payload: dict


class ExecuteResponse(BaseModel):
    results: dict
    errors: dict


# ---- planner (control_plane.py:45-75) ----------------------------------------------------------
def hash_embedding(text: str, dim: int = EMBED_DIM) -> np.ndarray:
    """Deterministic stand-in for the embedding model (no OpenAI offline): a standard-normal vector
    seeded by the intent's SHA-256."""
    seed = int.from_bytes(hashlib.sha256(text.encode("utf-8")).digest()[:8], "little")
    return np.random.default_rng(seed).standard_normal(dim).astype(np.float32)


def _openai_chat(prompt: str) -> str:
    import openai  # openai<1.0 API, as the reference uses it (:69-73)

    resp = openai.ChatCompletion.create(model="gpt-4o-mini", messages=[{"role": "system", "content": prompt}],
                                        temperature=0.2)
    return resp.choices[0].message.content


class MicroBatcher:
    """Coalesces concurrent single-query searches into one kernel launch.

    ``/plan`` runs on worker threads and ``/plan_and_execute`` on its own (control_plane.py:140-151), so several
    requests are usually in flight at once, yet each carries one intent; one scan of the table answers a whole batch
    in barely more time than a single query (the scan is memory-bound).  Callers block in ``search``; one worker
    thread drains whatever has queued up while the previous launch was running -- no timer, so a lone request pays
    no extra latency, and the batch size adapts to the load by itself.
    """

    def __init__(self, search_fn: Callable[[np.ndarray, int], tuple], max_batch: int = 1024):
        self._search, self._max = search_fn, max_batch
        self._cv = threading.Condition()
        self._pending: List[tuple] = []
        self._closed = False
        self.launches = self.requests = 0
        self._worker = threading.Thread(target=self._run, name="mcp-microbatch", daemon=True)
        self._worker.start()

    def search(self, query: np.ndarray, k: int):
        """One query vector -> (ids [k], scores [k]); blocks until its batch has been answered."""
        slot = {"q": np.asarray(query, np.float32), "k": int(k), "done": threading.Event(), "out": None, "err": None}
        with self._cv:
            if self._closed:
                raise RuntimeError("MicroBatcher is closed")
            self._pending.append(slot)
            self._cv.notify()
        slot["done"].wait()
        if slot["err"] is not None:
            raise slot["err"]
        return slot["out"]

    def _run(self):
        while True:
            with self._cv:
                while not self._pending and not self._closed:
                    self._cv.wait()
                if self._closed and not self._pending:
                    return
                batch, self._pending = self._pending[:self._max], self._pending[self._max:]
            try:
                kmax = max(s["k"] for s in batch)
                ids, scores, _ = self._search(np.stack([s["q"] for s in batch]), kmax)
                for i, s in enumerate(batch):
                    s["out"] = (ids[i, :s["k"]], scores[i, :s["k"]])     # a prefix of an exact top-kmax is the exact top-k
            except Exception as e:  # noqa: BLE001 -- handed to every waiting caller
                for s in batch:
                    s["err"] = e
            self.launches += 1
            self.requests += len(batch)
            for s in batch:
                s["done"].set()

    def close(self):
        with self._cv:
            self._closed = True
            self._cv.notify()
        self._worker.join(timeout=5)


class GraphPlanner:
    """Same constructor and ``plan`` signature as the reference; three optional injection points.

    fetch_embeddings: () -> [(name, vector)]   default: the reference's pgvector SELECT (:51-55)
    embed:            intent -> fp32[dim]      default: ``hash_embedding``
    llm:              prompt -> JSON text      default: the reference's ChatCompletion call
    """

    def __init__(self, registry: ServiceRegistry, pg_dsn: str = POSTGRES_DSN, *,
                 fetch_embeddings: Optional[Callable[[], Sequence]] = None,
                 embed: Optional[Callable[[str], np.ndarray]] = None,
                 llm: Optional[Callable[[str], str]] = None,
                 top_k: int = RETRIEVAL_TOP_K, metric: str = RETRIEVAL_METRIC, device=CUDA_DEVICE,
                 telemetry_client=None, micro_batch: bool = False):
        """``device``: a GPU ordinal, or a list of them -- the registry is then row-sharded over those GPUs behind the same
        ``plan`` / ``retrieve`` calls (``mcp_index_create_multi``)."""
        self.registry = registry
        self.pg_dsn = pg_dsn
        self.conn = None                       # opened lazily; the reference connects in __init__ (:48)
        self._fetch = fetch_embeddings
        self._embed = embed or hash_embedding
        self._llm = llm or _openai_chat
        self.top_k, self.metric, self.device = top_k, metric, device
        self._telemetry_client = telemetry_client
        self._index: Optional[ServiceIndex] = None
        self._names: List[str] = []
        self._lock = threading.Lock()
        self._batcher: Optional[MicroBatcher] = None
        self._micro_batch = micro_batch

    # the reference's only read of the embedding table
    def _fetch_embeddings_metadata(self):
        if self._fetch is not None:
            return self._fetch()
        import psycopg2
        from pgvector.psycopg2 import register_vector

        if self.conn is None:
            self.conn = psycopg2.connect(self.pg_dsn)
            register_vector(self.conn)
        with self.conn.cursor() as cur:
            cur.execute("SELECT name, input_schema_vector FROM service_schemas;")
            return cur.fetchall()

    # ---- table residency ----------------------------------------------------------------
    def refresh_index(self, services: Optional[List[dict]] = None) -> ServiceIndex:
        """(Re)load service_schemas into HBM and fold the telemetry penalty."""
        rows = self._fetch_embeddings_metadata()
        names = [r[0] for r in rows]
        table = np.stack([np.asarray(r[1], np.float32) for r in rows]) if rows else np.zeros((0, EMBED_DIM), np.float32)
        with self._lock:
            if self._index is None or self._index.dim != table.shape[1]:
                # a replaced index is not closed here: a search on another thread may still be inside it.  Searches hold their
                # own reference (retrieve_batch); the handle is destroyed when the last of them lets go (ServiceIndex.__del__)
                self._index = ServiceIndex(table.shape[1], self.metric, self.device)
                if self._batcher is not None:
                    self._batcher.close()                     # drains what is queued against the old index, then stops
                    self._batcher = None
            self._index.load(table)
            self._names = names
            if services is None:
                services = self.registry.list_services()
            by_name = {s["name"]: s for s in services}
            recs = [by_name.get(n, {"name": n}) for n in names]
            self._index.set_penalty(_telemetry.penalty_for_services(recs, self._telemetry_client))
        return self._index

    def attach_index(self, index: ServiceIndex, names: Sequence[str]) -> None:
        """Use an index that is already resident (e.g. adopted from device memory) instead of fetching the table."""
        with self._lock:
            self._index, self._names = index, list(names)

    def retrieve(self, intent: str, k: Optional[int] = None) -> List[tuple]:
        """Top-k (service name, score) for one intent.  With ``micro_batch=True`` concurrent callers share launches."""
        if not self._micro_batch:
            return self.retrieve_batch([intent], k)[0]
        if self._index is None:
            self.refresh_index()
        k = min(k or self.top_k, max(1, len(self._names)))
        with self._lock:
            if self._batcher is None:
                index = self._index
                self._batcher = MicroBatcher(lambda q, kk: index.search(q, kk))
            batcher, names = self._batcher, self._names
        ids, scores = batcher.search(self._embed(intent), k)
        return [(names[i], float(s)) for i, s in zip(ids, scores) if i >= 0]

    def retrieve_batch(self, intents: Sequence[str], k: Optional[int] = None) -> List[List[tuple]]:
        if self._index is None:
            self.refresh_index()
        with self._lock:
            index, names = self._index, self._names            # own references: a concurrent refresh may swap both
        k = min(k or self.top_k, max(1, len(names)))
        q = np.stack([np.asarray(self._embed(t), np.float32) for t in intents])
        ids, scores, _ = index.search(q, k)
        return [[(names[i], float(s)) for i, s in zip(row_i, row_s) if i >= 0]
                for row_i, row_s in zip(ids, scores)]

    # ---- plan (same contract as control_plane.py:57-75) -----------------------------------
    def plan(self, intent: str) -> PlanResponse:
        if self._index is None:
            self.refresh_index(self.registry.list_services())     # first request only: the reference scans every time (:58)
        names = [n for n, _ in self.retrieve(intent)]
        if hasattr(self.registry, "get_services"):
            candidates = self.registry.get_services(names)        # k GETs
        else:                                                     # a registry that only offers the reference's method
            by_name = {s["name"]: s for s in self.registry.list_services()}
            candidates = [by_name[n] for n in names if n in by_name]
        lines = ["You are an orchestration agent.  Given the user intent and available services,",
                 "output a JSON DAG specifying for each step: service_name, input_keys, next_steps, fallback.",
                 "", "Available services:"]
        lines += [f"- {s['name']} (endpoint: {s['endpoint']}, inputs: {s['input_schema']}, outputs: {s['output_schema']})"
                  for s in candidates]
        lines += ["", f"User intent: “{intent}”", "", "JSON DAG:"]
        return PlanResponse(graph=json.loads(self._llm("\n".join(lines))))


# ---- orchestrator (control_plane.py:87-131) ----------------------------------------------------
class Orchestrator:
    """``execute`` keeps the reference's order and semantics.  Two opt-in extras from SURVEY.md section 8f item 3:
    ``parallel_levels=True`` awaits every node of a generation together (the kernel hands back the generations; the
    reference awaits nodes one by one, :109), ``retries=n`` re-POSTs a failing service n more times before the
    fallback (README.md:49 promises retries, the reference has none)."""

    def __init__(self, client=None, device: int = CUDA_DEVICE, parallel_levels: bool = False, retries: int = 0):
        self.parallel_levels, self.retries = parallel_levels, retries
        self._init(client, device)

    def _init(self, client=None, device: int = CUDA_DEVICE):
        if client is None:
            import httpx

            client = httpx.AsyncClient()
        self.client = client
        self.device = device[0] if isinstance(device, (list, tuple)) else device   # level scheduling needs one GPU
        logging.basicConfig(level=logging.INFO)
        self.logger = logging.getLogger("orchestrator")

    # batched form: many candidate plan DAGs -> levels / order / status in one kernel launch
    def schedule_batch(self, graphs: Sequence[dict]):
        enc = _dag.encode_graphs(graphs)
        return enc, _dag.levels(enc, self.device, with_order=True)

    async def execute(self, graph: dict, payload: dict) -> ExecuteResponse:
        from fastapi import HTTPException
        import networkx as nx

        nodes: Dict[str, dict] = {}
        for node in graph["nodes"]:                       # first add_node fixes the position, later ones update attrs
            nodes.setdefault(node["name"], {}).update(node)
        first_in_edge: Dict[str, dict] = {}
        for edge in graph["edges"]:
            nodes.setdefault(edge["from"], {})
            nodes.setdefault(edge["to"], {})
            first_in_edge.setdefault(edge["to"], edge)
        # the ordering step (:104) on the GPU, off the event loop: the call blocks its thread until the kernel's answer is back
        enc, res = await asyncio.to_thread(self.schedule_batch, [graph])
        _, order, status = _dag.decode(enc, res, 0)

        results, errors = {}, {}

        async def run_node(name):
            node = nodes[name]
            service_url = node["endpoint"]                # KeyError for an undeclared node, as in the reference (:106)
            inputs = {k: results.get(v, payload.get(v)) for k, v in node["inputs"].items()}
            last = None
            for _ in range(1 + self.retries):
                try:
                    resp = await self.client.post(service_url, json=inputs, timeout=5.0)
                    resp.raise_for_status()
                    results[name] = resp.json()
                    return
                except Exception as e:  # noqa: BLE001 -- same breadth as the reference (:112)
                    last = e
            self.logger.error(f"Service {name} failed: {last}")
            errors[name] = str(last)
            fallback = (first_in_edge.get(name) or {}).get("fallback")
            if not fallback:
                raise HTTPException(status_code=502, detail=f"{name} failed and no fallback available")
            self.logger.info(f"Attempting fallback {fallback} for {name}")
            try:
                resp = await self.client.post(fallback, json=inputs, timeout=5.0)
                resp.raise_for_status()
                results[name] = resp.json()
            except Exception as e2:  # noqa: BLE001
                self.logger.error(f"Fallback {fallback} failed: {e2}")
                errors[name] += f"; fallback failed: {e2}"

        if self.parallel_levels:
            for gen in _dag.generations(enc, res, 0):     # nodes of one generation have no edges between them
                await asyncio.gather(*(run_node(n) for n in gen))
        else:
            for name in order:                            # == list(nx.topological_sort(G)), from the kernel
                await run_node(name)
        if status:
            # the reference raises only after the acyclic prefix has been POSTed
            raise nx.NetworkXUnfeasible("Graph contains a cycle or graph changed during iteration")
        return ExecuteResponse(results=results, errors=errors)


# ---- app (control_plane.py:135-151) -------------------------------------------------------------
def create_app(registry: Optional[ServiceRegistry] = None, planner: Optional[GraphPlanner] = None,
               orch: Optional[Orchestrator] = None):
    from fastapi import FastAPI

    app = FastAPI()
    registry = registry or ServiceRegistry(REDIS_URL)
    planner = planner or GraphPlanner(registry, POSTGRES_DSN)
    orch = orch or Orchestrator()
    app.state.registry, app.state.planner, app.state.orch = registry, planner, orch

    @app.post("/plan", response_model=PlanResponse)
    def plan_intent(req: PlanRequest):
        return planner.plan(req.intent)

    @app.post("/execute", response_model=ExecuteResponse)
    async def run_graph(req: ExecuteRequest):
        return await orch.execute(req.graph, req.payload)

    @app.post("/plan_and_execute", response_model=ExecuteResponse)
    async def plan_and_run(req: PlanRequest):
        plan = await asyncio.to_thread(planner.plan, req.intent)
        return await orch.execute(plan.graph, {})

    return app


if __name__ == "__main__":
    import uvicorn

    uvicorn.run(create_app(), host="0.0.0.0", port=8000)

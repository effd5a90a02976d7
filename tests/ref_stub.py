"""A stand-in for the un-vendored ``pyseekdb`` client so that the reference's UNMODIFIED connector runs here.

TEST INFRASTRUCTURE (used by tests/, tests/golden/make_golden.py and the C1 line of bench.py / tools/bench_c1.py).

``super_rag/vectorstore/seekdb_connector.py`` imports exactly three names from ``pyseekdb`` (``:1-2,7``): ``Client``,
``HNSWConfiguration`` and the generic ``EmbeddingFunction``. The real client talks to an external SeekDB / OceanBase
server (``deploy/docker-compose.yaml:2-3``) that cannot run in this sandbox. This module injects a ``pyseekdb`` module
into ``sys.modules`` whose collections live in host memory and whose ``query`` is the exact brute-force cosine oracle
(``oracle/vector.py`` arithmetic: fp32 matrix-vector product, row norms precomputed at ``add`` time, zero-norm -> 0,
order score desc / insertion row asc), returning the batched dict shape the connector unpacks (``:117-156``):

    {"ids": [[..]], "distances": [[..]], "documents": [[..]], "metadatas": [[..]]}

``distances`` carries the similarity (the connector forwards it verbatim as ``score``, ``:143,150``).
With it, ``SeekDBVectorStoreConnector.search`` (``:98-115``) and ``ContextManager.query`` (``context/context.py:14-48``)
execute unmodified — SURVEY §8(c) "How the reference connector itself is exercised for C1".

``install(canned=...)`` instead makes every query return a fixed dict (the golden generator pins the wrapping code that
way). Nothing here is reachable from the product package.
"""
from __future__ import annotations

import importlib
import os
import sys
import types
from typing import Any, Dict, List, Optional

import numpy as np

REFERENCE = os.environ.get("SRAG_REFERENCE", "/root/reference")


class _Generic:
    def __class_getitem__(cls, item):
        return cls


class StubCollection:
    """In-memory collection with the subset of the pyseekdb collection API the connector calls."""

    def __init__(self, name: str, state: "StubState"):
        self.name, self._state = name, state
        self.ids: List[str] = []
        self.documents: List[Optional[str]] = []
        self.metadatas: List[Optional[dict]] = []
        self._rows: List[np.ndarray] = []          # appended blocks; consolidated lazily
        self._matrix: Optional[np.ndarray] = None
        self._inv: Optional[np.ndarray] = None
        self._alive: Optional[np.ndarray] = None
        self.queries = 0

    # seekdb_connector.py:78-82
    def add(self, ids, embeddings=None, metadatas=None, documents=None):
        emb = np.asarray(embeddings, dtype=np.float32)
        if emb.ndim != 2 or emb.shape[0] != len(ids):
            raise ValueError("embeddings must be [len(ids), dim]")
        self.ids += list(ids)
        self.documents += list(documents) if documents is not None else [None] * len(ids)
        self.metadatas += list(metadatas) if metadatas is not None else [None] * len(ids)
        self._rows.append(emb)
        self._matrix = None

    # seekdb_connector.py:92-94
    def delete(self, ids):
        self._consolidate()
        dead = set(ids)
        for r, i in enumerate(self.ids):
            if i in dead:
                self._alive[r] = False

    def _consolidate(self):
        if self._matrix is None:
            old_alive = self._alive
            self._matrix = np.concatenate(self._rows) if self._rows else np.zeros((0, 1), np.float32)
            self._rows = [self._matrix]
            n = np.sqrt(np.einsum("ij,ij->i", self._matrix.astype(np.float64), self._matrix.astype(np.float64)))
            self._inv = np.where(n > 0, 1.0 / np.where(n > 0, n, 1.0), 0.0).astype(np.float32)
            self._alive = np.ones(self._matrix.shape[0], dtype=bool)
            if old_alive is not None:
                self._alive[:old_alive.shape[0]] = old_alive

    # seekdb_connector.py:103-107
    def query(self, query_embeddings=None, query_texts=None, n_results=None):
        self.queries += 1
        if self._state.canned is not None:
            return self._state.canned
        self._consolidate()
        q = np.asarray(query_embeddings, dtype=np.float32)
        k = int(n_results or 0)
        n = self._matrix.shape[0]
        if k <= 0 or n == 0:
            return {"ids": [[]], "distances": [[]], "documents": [[]], "metadatas": [[]]}
        nq = float(np.sqrt(np.dot(q.astype(np.float64), q.astype(np.float64))))
        qs = np.float32(1.0 / nq) if nq > 0 else np.float32(0.0)
        sc = (self._matrix @ q) * self._inv * qs                      # exact cosine, fp32 (oracle/vector.py arithmetic)
        sc = np.where(self._alive, sc, -np.inf)
        kk = min(k, int(self._alive.sum()))
        top = np.argpartition(-sc, kk - 1)[:kk] if kk < n else np.arange(n)
        kth = sc[top].min() if kk else np.inf
        cand = np.nonzero(sc >= kth)[0]                                # ties at the k-th: resolved by row below
        order = cand[np.lexsort((cand, -sc[cand].astype(np.float64)))][:kk]
        return {"ids": [[self.ids[r] for r in order]], "distances": [[float(sc[r]) for r in order]],
                "documents": [[self.documents[r] for r in order]], "metadatas": [[self.metadatas[r] for r in order]]}


class StubState:
    def __init__(self):
        self.collections: Dict[str, StubCollection] = {}
        self.canned: Optional[dict] = None


STATE = StubState()


class Client:
    """pyseekdb.Client(host=, port=, database=, user=, password=) — seekdb_connector.py:49-55."""

    def __init__(self, **kwargs: Any):
        self.kwargs = kwargs

    def create_collection(self, name, configuration=None, embedding_function=None):
        STATE.collections[name] = StubCollection(name, STATE)
        return STATE.collections[name]

    def get_or_create_collection(self, name):
        if name not in STATE.collections:
            STATE.collections[name] = StubCollection(name, STATE)
        return STATE.collections[name]

    def delete_collection(self, name):
        STATE.collections.pop(name, None)


def HNSWConfiguration(**kwargs):  # noqa: N802  (seekdb_connector.py:58-61; the stub is exact, the config is recorded only)
    return dict(kwargs)


def install(canned: Optional[dict] = None) -> types.ModuleType:
    """Put the stub into sys.modules as `pyseekdb` (idempotent) and return it."""
    STATE.canned = canned
    mod = sys.modules.get("pyseekdb")
    if mod is None or not getattr(mod, "_srag_stub", False):
        mod = types.ModuleType("pyseekdb")
        mod.Client, mod.HNSWConfiguration, mod.EmbeddingFunction = Client, HNSWConfiguration, _Generic
        mod._srag_stub = True
        sys.modules["pyseekdb"] = mod
    return mod


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE, "super_rag", "vectorstore", "seekdb_connector.py"))


def load_reference_connector():
    """-> (seekdb_connector module, ContextManager class, reference models module), all UNMODIFIED reference code imported
    from /root/reference with only `pyseekdb` stubbed. Raises FileNotFoundError where the reference tree is absent (the
    GPU box)."""
    if not reference_available():
        raise FileNotFoundError(f"{REFERENCE} is not present on this machine")
    install()
    if REFERENCE not in sys.path:
        sys.path.insert(0, REFERENCE)
    seek = importlib.import_module("super_rag.vectorstore.seekdb_connector")
    ctx = importlib.import_module("super_rag.context.context")
    models = importlib.import_module("super_rag.models")
    return seek, ctx.ContextManager, models

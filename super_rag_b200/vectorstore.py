"""Drop-in vector-store connector: same duck type as the reference's
``SeekDBVectorStoreConnector`` (super_rag/vectorstore/seekdb_connector.py:31-156), with the
``collection.query`` round-trip to an external HNSW server replaced by the exact B200 scan (K1).

Selected with ``VECTOR_DB_TYPE=b200`` through ``VectorStoreConnectorAdaptor``
(super_rag/vectorstore/connector.py:4-15 gains ``case "b200"``). The reference
constructs a fresh connector per query (nodeflow/runners/vector_search.py:74 ->
context/context.py:12), so device state lives in a process-global registry keyed by collection
name, not in the connector instance.

Scores are similarities (higher is better): cosine or inner product, order (score desc, chunk
row asc). ``score_threshold`` / ``filter`` kwargs are accepted and ignored exactly as the
reference does (seekdb_connector.py:99-100) unless ctx["apply_score_threshold"] / ctx["apply_filter"]
is set; with ``apply_filter`` the metadata filter built by ``ContextManager._create_combined_filter``
(context/context.py:74-111) is folded into the scan.
"""
from __future__ import annotations

import logging
import threading
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops, shardio
from ._lib import SragError
from .doctable import HostDocTable
from .models import DocumentWithScore, QueryResult, QueryWithEmbedding

logger = logging.getLogger(__name__)


MISSING_BIT = 1 << 31      # row has no `indexer` field ({"indexer": {"$exists": False}})
NO_SUCH_GROUP = 0x7FFFFFFF  # chat_id never seen in this collection: matches no row


def _i32(pattern: int) -> int:
    """32-bit pattern as the signed value an int32 tensor stores."""
    pattern &= 0xFFFFFFFF
    return pattern - (1 << 32) if pattern & 0x80000000 else pattern


class RowTagger:
    """Interns the two metadata fields the reference filters on (context/context.py:74-111): `indexer`
    -> one bit of a 32-bit tag word (bit 31 = field absent), `chat_id` -> a small integer (0 = absent)."""

    def __init__(self):
        self.indexer_bit: Dict[str, int] = {}
        self.chat_group: Dict[str, int] = {}

    def encode(self, meta: Optional[dict]) -> Tuple[int, int]:
        meta = meta or {}
        tag = MISSING_BIT
        if meta.get("indexer") is not None:
            name = str(meta["indexer"])
            if name not in self.indexer_bit:
                if len(self.indexer_bit) >= 31:
                    raise ValueError("more than 31 distinct `indexer` values in one collection")
                self.indexer_bit[name] = len(self.indexer_bit)
            tag = 1 << self.indexer_bit[name]
        group = 0
        if meta.get("chat_id") is not None:
            group = self.chat_group.setdefault(str(meta["chat_id"]), len(self.chat_group) + 1)
        return tag, group

    def encode_batch(self, metas: Sequence[Optional[dict]]) -> Tuple[np.ndarray, np.ndarray]:
        """-> (tag words, chat groups) as int32 arrays for a batch of metadata dicts. Rows sharing one dict object
        (the chunks of one document usually do) are encoded once; rows without metadata never enter Python code
        beyond the loop header."""
        m = len(metas)
        tags = [MISSING_BIT] * m
        groups = [0] * m
        last, last_enc = None, (MISSING_BIT, 0)
        for j, mt in enumerate(metas):
            if not mt:
                continue
            if mt is not last:
                last, last_enc = mt, self.encode(mt)
            tags[j], groups[j] = last_enc
        return (np.array(tags, dtype=np.uint32).view(np.int32), np.array(groups, dtype=np.int32))

    def state(self) -> dict:
        return {"indexer_bit": self.indexer_bit, "chat_group": self.chat_group}

    @classmethod
    def from_state(cls, st: dict) -> "RowTagger":
        t = cls()
        t.indexer_bit = {str(k): int(v) for k, v in st.get("indexer_bit", {}).items()}
        t.chat_group = {str(k): int(v) for k, v in st.get("chat_group", {}).items()}
        return t

    # -- filter dict -> (tag mask, group) ------------------------------------------------------
    def _tag_clause(self, cond) -> int:
        if isinstance(cond, dict):
            if set(cond) == {"$in"}:
                m = 0
                for v in cond["$in"]:
                    b = self.indexer_bit.get(str(v))
                    m |= 0 if b is None else 1 << b
                return m
            if set(cond) == {"$exists"}:
                return MISSING_BIT if not cond["$exists"] else (0xFFFFFFFF & ~MISSING_BIT)
            raise ValueError(f"unsupported filter operator {cond!r}")
        b = self.indexer_bit.get(str(cond))
        return 0 if b is None else 1 << b

    def compile(self, flt) -> Tuple[int, int]:
        """-> (tag mask, group); mask 0xFFFFFFFF / group -1 mean "no constraint". Grammar = what
        `_create_combined_filter` emits, plus bare `indexer` clauses: {"or": [tag clauses]}, {"chat_id": x},
        {"indexer": v | {"$in": [...]} | {"$exists": bool}}, {"and": [clauses]}."""
        if not flt:
            return 0xFFFFFFFF, -1
        if not isinstance(flt, dict) or len(flt) != 1:
            raise ValueError(f"unsupported filter {flt!r}")
        (key, val), = flt.items()
        if key == "and":
            mask, group = 0xFFFFFFFF, -1
            for c in val:
                m, g = self.compile(c)
                mask &= m
                if g >= 0:
                    if group >= 0 and g != group:
                        group = NO_SUCH_GROUP
                    else:
                        group = g
            return mask, group
        if key == "or":
            mask = 0
            for c in val:
                m, g = self.compile(c)
                if g >= 0:
                    raise ValueError("`chat_id` inside an `or` is not supported")
                mask |= m
            return mask, -1
        if key == "indexer":
            return self._tag_clause(val), -1
        if key == "chat_id":
            return 0xFFFFFFFF, self.chat_group.get(str(val), NO_SUCH_GROUP)
        raise ValueError(f"unsupported filter field {key!r}")


class DeviceCollection:
    """bf16 corpus matrix + fp32 row multipliers resident in HBM, growable, with tombstones."""

    def __init__(self, name: str, dim: int, metric="cosine", device: Optional[str] = None, capacity: int = 0):
        if dim % 8 != 0 or dim < 8:
            raise ValueError(f"vector_size={dim} must be a positive multiple of 8")
        self.name, self.dim, self.metric = name, dim, ops.metric_code(metric)
        self.device = torch.device(device) if device else None
        self.docs = HostDocTable()
        self.n = 0
        self.id_base = 0
        self._rows: Optional[torch.Tensor] = None
        self._scale: Optional[torch.Tensor] = None
        self._ws: Optional[torch.Tensor] = None
        self._ws_event: Optional[torch.cuda.Event] = None   # last search that used the workspace
        self._tags: Optional[torch.Tensor] = None      # int32 [cap] tag words, int32 [cap] chat groups
        self._group: Optional[torch.Tensor] = None
        self.tagger = RowTagger()
        self._reserve = capacity
        self.lock = threading.Lock()

    # -- storage ------------------------------------------------------------------------------
    def _dev(self) -> torch.device:
        if self.device is None:
            if not torch.cuda.is_available():
                raise SragError("B200VectorStoreConnector needs a CUDA device: there is no CPU fallback")
            self.device = torch.device("cuda", torch.cuda.current_device())
        return self.device

    def _grow(self, need: int):
        cap = 0 if self._rows is None else self._rows.shape[0]
        if need <= cap:
            return
        new_cap = max(need, 2 * cap, self._reserve, 1024)
        dev = self._dev()
        rows = torch.empty(new_cap, self.dim, dtype=torch.bfloat16, device=dev)
        scale = torch.empty(new_cap, dtype=torch.float32, device=dev)
        tags = torch.empty(new_cap, dtype=torch.int32, device=dev)
        group = torch.empty(new_cap, dtype=torch.int32, device=dev)
        if self.n:
            rows[:self.n].copy_(self._rows[:self.n])
            scale[:self.n].copy_(self._scale[:self.n])
            tags[:self.n].copy_(self._tags[:self.n])
            group[:self.n].copy_(self._group[:self.n])
        self._rows, self._scale, self._tags, self._group = rows, scale, tags, group

    @property
    def rows(self) -> torch.Tensor:
        return self._rows[:self.n] if self._rows is not None else torch.empty(0, self.dim, dtype=torch.bfloat16,
                                                                              device=self._dev())

    @property
    def row_scale(self) -> torch.Tensor:
        return self._scale[:self.n] if self._scale is not None else torch.empty(0, dtype=torch.float32,
                                                                                device=self._dev())

    def append(self, embeddings, ids=None, texts=None, metas=None) -> List[str]:
        """embeddings: [m, dim] tensor (any device, fp32 or bf16) or nested lists. All-or-nothing: lengths and
        duplicate ids are rejected before the device or the doc table is touched."""
        emb = embeddings if isinstance(embeddings, torch.Tensor) else torch.tensor(embeddings, dtype=torch.float32)
        if emb.ndim != 2 or emb.shape[1] != self.dim:
            raise ValueError(f"embeddings must be [m, {self.dim}], got {tuple(emb.shape)}")
        m = emb.shape[0]
        ids = list(ids) if ids is not None else None           # None: uuid4s generated in bulk (seekdb_connector.py:75)
        texts = list(texts) if texts is not None else None
        metas = list(metas) if metas is not None else None
        with self.lock:
            self.docs.validate_new(ids, m, texts, metas)
            tagger_before = (dict(self.tagger.indexer_bit), dict(self.tagger.chat_group))
            try:
                tags, groups = self.tagger.encode_batch(metas) if metas is not None else (
                    np.full(m, _i32(MISSING_BIT), dtype=np.int32), np.zeros(m, dtype=np.int32))
            except Exception:
                self.tagger.indexer_bit, self.tagger.chat_group = tagger_before
                raise
            dev = self._dev()
            self._grow(self.n + m)
            if emb.dtype not in (torch.float32, torch.bfloat16):
                emb = emb.to(torch.float32)
            emb = emb.to(dev, non_blocking=True).contiguous()
            ops.ingest_rows(emb, self.metric, out_bf16=self._rows[self.n:self.n + m],
                            out_scale=self._scale[self.n:self.n + m])
            self._tags[self.n:self.n + m] = torch.from_numpy(tags).to(dev)
            self._group[self.n:self.n + m] = torch.from_numpy(groups).to(dev)
            out_ids = self.docs.append(ids, texts, metas, m)    # cannot fail any more: validated above
            self.n += m
        return out_ids

    def delete(self, ids: Sequence[str]) -> int:
        with self.lock:
            rows = self.docs.delete(ids)
            if rows:
                idx = torch.tensor(rows, dtype=torch.int64, device=self._dev())
                self._scale[idx] = float("nan")  # tombstone: K1 never returns a NaN-scaled row
        return len(rows)

    # -- persistence (on-disk shard) ---------------------------------------
    def save(self, path: str) -> None:
        """Write rows, multipliers (NaN tombstones included), the filter columns and the doc table (id / text / metadata
        columns as raw sections) as one shard file."""
        with self.lock:
            meta = {"name": self.name, "dim": self.dim, "metric": self.metric, "n": self.n, "id_base": self.id_base,
                    "doc_table": "columns"}
            arrays = {"rows": self.rows, "row_scale": self.row_scale}
            if self.n:
                arrays["row_tags"], arrays["row_group"] = self._tags[:self.n], self._group[:self.n]
            arrays.update({k: torch.from_numpy(np.array(v)) for k, v in self.docs.to_arrays().items()})
            shardio.write_shard(path, "vector", meta, arrays, {"tagger": self.tagger.state()})

    @classmethod
    def load(cls, path: str, name: Optional[str] = None, device: Optional[str] = None) -> "DeviceCollection":
        """Rebuild a collection from a shard written by save(); searches return bit-identical results. The doc table
        comes back as columns (no per-row Python work; metadata is parsed when a hit asks for it)."""
        footer = shardio.read_footer(path)
        meta = footer["meta"]
        col = cls(name or meta["name"], int(meta["dim"]), int(meta["metric"]), device)
        dev = col._dev()
        host_sections = lambda nm: nm.startswith("doc_")
        meta, arrays, blobs = shardio.read_shard(path, device=dev, expect_kind="vector", host_sections=host_sections)
        n = int(meta["n"])
        rows, scale = arrays["rows"], arrays["row_scale"]
        if tuple(rows.shape) != (n, col.dim) or rows.dtype != torch.bfloat16 or tuple(scale.shape) != (n,):
            raise shardio.ShardError(f"{path}: rows / row_scale do not match the header")
        col._rows, col._scale, col.n, col.id_base = rows, scale, n, int(meta["id_base"])
        if meta.get("doc_table") == "columns":
            try:
                col.docs = HostDocTable.from_arrays({k: v.numpy() for k, v in arrays.items() if k.startswith("doc_")}, n)
            except (KeyError, ValueError) as e:
                raise shardio.ShardError(f"{path}: doc table does not match the header ({e})") from e
            col.tagger = RowTagger.from_state(blobs.get("tagger", {}))
            if n:
                col._tags, col._group = arrays["row_tags"].contiguous(), arrays["row_group"].contiguous()
                if col._tags.numel() != n or col._group.numel() != n:
                    raise shardio.ShardError(f"{path}: filter columns do not match the header")
            else:
                col._tags = torch.empty(0, dtype=torch.int32, device=dev)
                col._group = torch.empty(0, dtype=torch.int32, device=dev)
            return col
        d = blobs["docs"]                                     # shards written before the column format
        if not (len(d["ids"]) == len(d["texts"]) == len(d["metas"]) == n):
            raise shardio.ShardError(f"{path}: doc table does not match the header")
        col.docs = HostDocTable.from_lists(d["ids"], d["texts"], d["metas"], d["deleted"])
        tags, groups = col.tagger.encode_batch(list(d["metas"]))   # same interning order as the original appends
        col._tags, col._group = torch.from_numpy(tags).to(dev), torch.from_numpy(groups).to(dev)
        return col

    # -- search -------------------------------------------------------------------------------
    def search_batch(self, queries: torch.Tensor, top_k: int, filters=None):
        """queries [B, dim] (host or device, fp32 or bf16) -> (scores fp32 [B,k], ids int64 [B,k]) on device.
        `filters`: None, one filter dict for the whole batch, or a list of B dicts (context.py:74-111 grammar);
        the result is the exact top-k among the rows each query's filter admits.

        The collection lock is held from the snapshot of (n, rows, multipliers, filter columns) to the last launch:
        the scan workspace (per-thread heaps, shared thresholds, pacing counters) belongs to the collection, ctypes
        releases the GIL during srag_* calls, and two interleaved searches — or a search racing an append that swaps
        the buffers — would silently merge each other's state."""
        dev = self._dev()
        q = queries.to(dev, non_blocking=True)
        if q.dtype not in (torch.float32, torch.bfloat16):
            q = q.to(torch.float32)
        with self.lock:
            n = self.n
            rows, scale = self.rows, self.row_scale
            qb, qs = ops.ingest_rows(q.contiguous(), self.metric)
            need = ops.topk_workspace_bytes(n, min(q.shape[0], ops.MAX_QUERIES_PER_CALL), top_k, self.dim)
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
            kw = {}
            if filters is not None and n:
                per_q = filters if isinstance(filters, (list, tuple)) else [filters] * q.shape[0]
                if len(per_q) != q.shape[0]:
                    raise ValueError(f"{len(per_q)} filters for {q.shape[0]} queries")
                comp = [self.tagger.compile(f) for f in per_q]
                if any(m != 0xFFFFFFFF for m, _ in comp):
                    kw["row_tags"] = self._tags[:n]
                    kw["q_tag_mask"] = torch.tensor([_i32(m) for m, _ in comp], dtype=torch.int32).to(dev)
                if any(g >= 0 for _, g in comp):
                    kw["row_group"] = self._group[:n]
                    kw["q_group"] = torch.tensor([g for _, g in comp], dtype=torch.int32).to(dev)
            if self._ws_event is not None:   # a search issued on another stream may still be using the workspace
                torch.cuda.current_stream(dev).wait_event(self._ws_event)
            out = ops.topk_search(rows, scale, qb, qs, top_k, id_base=self.id_base, workspace=self._ws, **kw)
            self._ws_event = torch.cuda.Event()
            self._ws_event.record(torch.cuda.current_stream(dev))
        return out


_REGISTRY: Dict[str, DeviceCollection] = {}
_REGISTRY_LOCK = threading.Lock()


class B200Client:
    """What ``connector.client`` exposes (the reference hands out its pyseekdb client there,
    seekdb_connector.py:49-55): collection lifecycle in the process-global registry."""

    def create_collection(self, name: str, dim: int, metric="cosine", device=None, capacity: int = 0):
        with _REGISTRY_LOCK:
            if name in _REGISTRY:
                raise ValueError(f"collection {name!r} already exists")
            _REGISTRY[name] = DeviceCollection(name, dim, metric, device, capacity)
            return _REGISTRY[name]

    def get_or_create_collection(self, name: str, dim: int = 1024, metric="cosine", device=None):
        with _REGISTRY_LOCK:
            if name not in _REGISTRY:
                _REGISTRY[name] = DeviceCollection(name, dim, metric, device)
            return _REGISTRY[name]

    def get_collection(self, name: str) -> DeviceCollection:
        return _REGISTRY[name]

    def has_collection(self, name: str) -> bool:
        return name in _REGISTRY

    def delete_collection(self, name: str):
        with _REGISTRY_LOCK:
            _REGISTRY.pop(name, None)

    def save_collection(self, name: str, path: str) -> None:
        _REGISTRY[name].save(path)

    def load_collection(self, path: str, name: Optional[str] = None, device=None) -> DeviceCollection:
        """Register the collection stored at `path` (replacing one of the same name)."""
        col = DeviceCollection.load(path, name, device)
        with _REGISTRY_LOCK:
            _REGISTRY[col.name] = col
        return col


class B200VectorStoreConnector:
    """Same constructor, attributes and methods as SeekDBVectorStoreConnector."""

    def __init__(self, ctx: Dict[str, Any], **kwargs: Any):
        self.ctx = ctx
        self.vector_size = ctx.get("vector_size", 1024)      # seekdb_connector.py:37
        self.distance = ctx.get("distance", "cosine")        # seekdb_connector.py:38
        self.collection_name = self.ctx["collection"]        # KeyError like the reference
        self.device = ctx.get("device")
        self.apply_score_threshold = bool(ctx.get("apply_score_threshold", False))
        self.apply_filter = bool(ctx.get("apply_filter", False))
        self.store = self                                    # seekdb_connector.py:45
        self.client = B200Client()

    def _collection(self) -> DeviceCollection:
        return self.client.get_or_create_collection(self.collection_name, self.vector_size, self.distance, self.device)

    def create_collection(self, **kwargs: Any):
        vector_size = kwargs.get("vector_size") or self.vector_size
        self.vector_size = vector_size
        if not self.client.has_collection(self.collection_name):
            self.client.create_collection(self.collection_name, vector_size, self.distance, self.device)

    def add(self, nodes) -> List[str]:
        """nodes: objects with .text/.metadata/.embedding (models.TextNode); returns new uuid4 ids."""
        nodes = list(nodes)
        if not nodes:
            return []
        emb = [n.embedding for n in nodes]
        if isinstance(emb[0], torch.Tensor):
            emb = torch.stack([e.to(torch.float32) for e in emb])
        return self._collection().append(emb, texts=[n.text for n in nodes], metas=[n.metadata for n in nodes])

    def delete_collection(self):
        self.client.delete_collection(self.collection_name)

    def delete(self, **delete_kwargs: Any):
        ids = delete_kwargs.get("ids")
        if ids:
            self._collection().delete(ids)
        else:
            raise ValueError("ids is required")              # seekdb_connector.py:95-96

    def search(self, query: QueryWithEmbedding, **kwargs) -> QueryResult:
        score_threshold = kwargs.get("score_threshold", 0.1)
        col = self._collection()
        top_k = int(query.top_k or 0)
        results: List[DocumentWithScore] = []
        if top_k > 0 and col.n > 0:
            k = min(top_k, ops.MAX_TOPK)
            q = torch.tensor([query.embedding], dtype=torch.float32)
            scores, rows = col.search_batch(q, k, filters=kwargs.get("filter") if self.apply_filter else None)
            scores, rows = scores[0].tolist(), rows[0].tolist()
            for s, r in zip(scores, rows):
                if r < 0:
                    break
                if self.apply_score_threshold and s < score_threshold:
                    break
                _id, text, meta = col.docs.get(r - col.id_base)
                results.append(DocumentWithScore(text=text, metadata=meta, score=s))
        return QueryResult(query=query.query, results=results)

    # batched tensor API used by benchmarks and the sharded index
    def search_batch(self, queries: torch.Tensor, top_k: int, filters=None):
        return self._collection().search_batch(queries, top_k, filters)


_REFERENCE_ADAPTOR = None  # set by nodeflow_pack.install_connector() to the class it replaces


class VectorStoreConnectorAdaptor:
    """super_rag/vectorstore/connector.py:4-15 with the extra ``case "b200"``; any other type is
    delegated to the reference adaptor when it is importable."""

    _srag_b200 = True

    def __init__(self, vector_store_type, ctx: Dict[str, Any], **kwargs: Any) -> None:
        self.ctx = ctx
        self.vector_store_type = vector_store_type
        match vector_store_type:
            case "b200":
                self.connector = B200VectorStoreConnector(ctx, **kwargs)
            case _:
                ref = _REFERENCE_ADAPTOR
                if ref is None:
                    try:
                        from super_rag.vectorstore.connector import VectorStoreConnectorAdaptor as ref  # type: ignore
                    except Exception:
                        ref = None
                if ref is None or getattr(ref, "_srag_b200", False):
                    raise ValueError("unsupported vector store type:", vector_store_type)  # connector.py:14-15
                self.connector = ref(vector_store_type, ctx, **kwargs).connector

"""Nodeflow node pack: ``fulltext_search`` and an extended ``merge`` (plus a stand-alone
``vector_search``), registered through the reference's own plugin mechanism.

Reference contract (paths relative to the reference root):
  * ``@register_node_runner(node_type, input_model=, output_model=)`` stores
    ``{"runner": cls(), "input_model", "output_model"}`` in ``NODE_RUNNER_REGISTRY``
    (super_rag/nodeflow/base/models.py:122-144); re-registering a type overwrites it (:137).
  * runner signature ``async def run(self, ui, si: SystemInput) -> (output_model, dict)`` (:125-128).
  * packs are zero-argument callables in entry-point group ``super_rag.nodeflow.packs``
    (super_rag/nodeflow/registry.py:19,44-60); this module's ``register`` is that callable
    (pyproject.toml of this repo declares it).

Inside a super-rag checkout the reference's registry, ``BaseNodeRunner``, ``SystemInput`` and
``ValidationError`` are used; stand-alone, field-for-field mirrors below keep the nodes usable and
testable. Node arithmetic (BM25 scoring, RRF / weighted fusion) runs on the GPU through the C ABI;
``union`` is the reference's host code, statement for statement in behaviour.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Callable, Dict, List, Optional, Tuple

from pydantic import BaseModel, Field

from .models import DocumentWithScore

try:  # pragma: no cover - exercised inside a super-rag checkout
    from super_rag.nodeflow.base.exceptions import ValidationError  # type: ignore
    from super_rag.nodeflow.base.models import (  # type: ignore
        NODE_RUNNER_REGISTRY, BaseNodeRunner, SystemInput, register_node_runner)
    USING_REFERENCE_NODEFLOW = True
except Exception:
    USING_REFERENCE_NODEFLOW = False
    NODE_RUNNER_REGISTRY: Dict[str, Dict[str, Any]] = {}

    class ValidationError(Exception):  # nodeflow/base/exceptions.py: ValidationError(nodeflowError)
        """Raised when input validation fails"""

    class BaseNodeRunner(ABC):  # nodeflow/base/models.py:125-128
        @abstractmethod
        async def run(self, ui: Any, si: Any) -> Tuple[Any, Dict[str, Any]]:
            raise NotImplementedError

    def register_node_runner(node_type: str, input_model, output_model):  # nodeflow/base/models.py:131-144
        def decorator(cls):
            NODE_RUNNER_REGISTRY[node_type] = {"runner": cls(), "input_model": input_model,
                                               "output_model": output_model}
            return cls
        return decorator

    class SystemInput:  # nodeflow/base/models.py:147-168
        def __init__(self, query: str, user: str, history=None, message_id: Optional[str] = None, **kwargs):
            self.query, self.user, self.history, self.message_id = query, user, history, message_id
            self.chat_id = None
            for key, value in kwargs.items():
                setattr(self, key, value)


# ---------------------------------------------------------------------------------------------
# fulltext_search
# ---------------------------------------------------------------------------------------------
class FulltextSearchInput(BaseModel):
    """FulltextSearchParams{topk, keywords} (schema/view_models.py:1043-1047) in the shape of
    VectorSearchInput (runners/vector_search.py:24-28)."""
    top_k: int = Field(5, description="Number of top results to return")
    keywords: Optional[List[str]] = Field(None, description="Custom keywords to use for fulltext search")
    collection_ids: Optional[List[str]] = Field(default_factory=list, description="Collection IDs")
    chat_id: Optional[str] = Field(None, description="Chat ID to filter chat documents")


class FulltextSearchOutput(BaseModel):
    docs: List[DocumentWithScore]


class FulltextSearchNodeRunner(BaseNodeRunner):
    """BM25 recall over the collection's chunk texts (K2). Mirrors VectorSearchNodeRunner.run
    (runners/vector_search.py:118-135): first collection id only, ``recall_type`` tag, any failure
    degrades to an empty list (:95-105)."""

    async def run(self, ui: FulltextSearchInput, si: SystemInput) -> Tuple[FulltextSearchOutput, dict]:
        import logging
        from . import fulltext
        collection_ids = ui.collection_ids or getattr(si, "collection_ids", []) or []
        docs: List[DocumentWithScore] = []
        if collection_ids:
            col = fulltext.get_collection(str(collection_ids[0]))  # generate_fulltext_index_name: str(id)
            if col is not None:
                try:
                    # chat documents stay inside their chat, as the vector node asks of its store
                    # (runners/vector_search.py:79-86 -> context.py:74-111); honoured by collections created with
                    # apply_filter=True
                    flt = {"chat_id": ui.chat_id} if getattr(ui, "chat_id", None) else None
                    docs = col.search(getattr(si, "query", None), ui.top_k, ui.keywords, filter=flt)
                    for item in docs:
                        if item.metadata is None:
                            item.metadata = {}
                        item.metadata["recall_type"] = "fulltext_search"
                except Exception as e:  # same degradation as the vector node
                    logging.getLogger(__name__).error(f"Fulltext search failed for collection {col.name}: {e}")
                    docs = []
        return FulltextSearchOutput(docs=docs), {}


# ---------------------------------------------------------------------------------------------
# merge (union | rrf | weighted)
# ---------------------------------------------------------------------------------------------
class MergeInput(BaseModel):
    """Reference MergeInput (runners/merge.py:12-28) + the fields the new strategies need."""
    merge_strategy: str = Field("union", description="How to merge results: union | rrf | weighted")
    deduplicate: bool = Field(True, description="Whether to deduplicate merged results")
    vector_search_docs: Optional[List[DocumentWithScore]] = Field(default_factory=list, description="Vector search docs")
    fulltext_search_docs: Optional[List[DocumentWithScore]] = Field(default_factory=list,
                                                                  description="Fulltext search docs")
    graph_search_docs: Optional[List[DocumentWithScore]] = Field(default_factory=list, description="Graph search docs")
    summary_search_docs: Optional[List[DocumentWithScore]] = Field(default_factory=list,
                                                                 description="Summary search docs")
    vision_search_docs: Optional[List[DocumentWithScore]] = Field(default_factory=list, description="Vision search docs")
    top_k: Optional[int] = Field(None, description="Keep only the best top_k fused results (rrf / weighted)")
    rank_const: int = Field(1, description="RRF rank constant (graphiti rrf default 1)")
    min_score: float = Field(0.0, description="Drop fused results scoring below this")
    weights: Optional[List[float]] = Field(None, description="Per-list weights for 'weighted' (vector, fulltext, "
                                                             "graph, summary, vision)")


class MergeOutput(BaseModel):
    docs: List[DocumentWithScore]


def fuse_documents(lists: List[List[DocumentWithScore]], strategy: str, top_k: Optional[int] = None,
                   rank_const: int = 1, min_score: float = 0.0, weights: Optional[List[float]] = None
                   ) -> List[DocumentWithScore]:
    """RRF / weighted fusion of document lists on the GPU (K3).

    Documents carry no id (models.py:263-266), so — like the reference's union — identity is the
    text: each distinct text gets a dense id in first-seen order over (list, position). With that id,
    the kernel's (score desc, id asc) order equals the reference rrf's stable first-seen tie order
    (graphiti_core/search/search_utils.py:1763-1778).
    """
    import torch
    from . import ops
    lists = [l for l in lists]
    id_of: Dict[Optional[str], int] = {}
    first_doc: List[DocumentWithScore] = []
    width = max([len(l) for l in lists] + [1])
    ids = torch.full((len(lists), 1, width), -1, dtype=torch.int64)
    scores = torch.zeros((len(lists), 1, width), dtype=torch.float32)
    for li, l in enumerate(lists):
        for pi, doc in enumerate(l):
            key = doc.text
            if key not in id_of:
                id_of[key] = len(first_doc)
                first_doc.append(doc)
            ids[li, 0, pi] = id_of[key]
            scores[li, 0, pi] = float(doc.score) if doc.score is not None else 0.0  # rerank.py:193: None -> 0.0
    n = len(first_doc)
    if n == 0:
        return []
    k_out = n if top_k is None else max(0, min(int(top_k), n))
    if k_out == 0:
        return []
    if len(lists) * width > 2048 or len(lists) > ops.MAX_LISTS:
        raise ValidationError(f"too many documents to fuse: {len(lists)} lists x {width}")
    dev = torch.device("cuda", torch.cuda.current_device())
    mode = ops.FUSE_RRF if strategy == "rrf" else ops.FUSE_WEIGHTED
    w = None
    if mode == ops.FUSE_WEIGHTED:
        w = list(weights) if weights else [1.0] * len(lists)
        if len(w) != len(lists):
            raise ValidationError(f"weights must have {len(lists)} entries, got {len(w)}")
    out_ids, out_scores = ops.fuse(ids.to(dev), scores.to(dev), mode, k_out, rank_const=rank_const,
                                   min_score=min_score, weights=w)
    docs: List[DocumentWithScore] = []
    for i, s in zip(out_ids[0].tolist(), out_scores[0].tolist()):
        if i < 0:
            break
        src = first_doc[i]
        docs.append(DocumentWithScore(text=src.text, metadata=src.metadata, score=s))
    return docs


class MergeNodeRunner(BaseNodeRunner):
    async def run(self, ui: MergeInput, si: SystemInput) -> Tuple[MergeOutput, dict]:
        docs_a: List[DocumentWithScore] = ui.vector_search_docs or []
        docs_b: List[DocumentWithScore] = ui.fulltext_search_docs or []
        docs_c: List[DocumentWithScore] = ui.graph_search_docs or []
        docs_d: List[DocumentWithScore] = ui.summary_search_docs or []
        docs_e: List[DocumentWithScore] = ui.vision_search_docs or []
        merge_strategy: str = ui.merge_strategy
        deduplicate: bool = ui.deduplicate

        if merge_strategy not in ["union", "rrf", "weighted"]:
            raise ValidationError(f"Unknown merge strategy: {merge_strategy}")

        if merge_strategy == "union":  # runners/merge.py:56-65, unchanged behaviour
            all_docs = docs_a + docs_b + docs_c + docs_d + docs_e
            if deduplicate:
                seen = set()
                unique_docs = []
                for doc in all_docs:
                    if doc.text not in seen:
                        seen.add(doc.text)
                        unique_docs.append(doc)
                return MergeOutput(docs=unique_docs), {}
            return MergeOutput(docs=all_docs), {}

        lists = [docs_a, docs_b, docs_c, docs_d, docs_e]
        weights = ui.weights
        if weights is not None and len(weights) != len(lists):
            raise ValidationError(f"weights must have {len(lists)} entries, got {len(weights)}")
        docs = fuse_documents(lists, merge_strategy, ui.top_k, ui.rank_const, ui.min_score, weights)
        return MergeOutput(docs=docs), {}


# ---------------------------------------------------------------------------------------------
# stand-alone vector_search (only registered when the reference's own runner is absent)
# ---------------------------------------------------------------------------------------------
class VectorSearchInput(BaseModel):  # runners/vector_search.py:24-28
    top_k: int = Field(5, description="Number of top results to return")
    similarity_threshold: float = Field(0.2, description="Similarity threshold for vector search")
    collection_ids: Optional[List[str]] = Field(default_factory=list, description="Collection IDs")
    chat_id: Optional[str] = Field(None, description="Chat ID to filter chat documents")


class VectorSearchOutput(BaseModel):  # runners/vector_search.py:32-33
    docs: List[DocumentWithScore]


_EMBEDDER: Optional[Callable[[str], List[float]]] = None


def set_embedder(fn: Optional[Callable[[str], List[float]]]) -> None:
    """Stand-alone only: text -> embedding (the reference resolves the collection's embedding
    service, runners/vector_search.py:70-76)."""
    global _EMBEDDER
    _EMBEDDER = fn


class B200VectorSearchNodeRunner(BaseNodeRunner):
    """VectorSearchNodeRunner.run / execute_vector_search (runners/vector_search.py:52-135) without
    the database and provider lookups: the collection name is the collection id, the query vector
    comes from ``si.query_vector`` or the registered embedder."""

    async def run(self, ui: VectorSearchInput, si: SystemInput) -> Tuple[VectorSearchOutput, dict]:
        import logging
        from .models import QueryWithEmbedding
        from .vectorstore import VectorStoreConnectorAdaptor
        collection_ids = ui.collection_ids or getattr(si, "collection_ids", []) or []
        if not collection_ids:
            return VectorSearchOutput(docs=[]), {}
        try:
            vector = getattr(si, "query_vector", None)
            if vector is None:
                if _EMBEDDER is None:
                    raise RuntimeError("no query_vector in SystemInput and no embedder registered")
                vector = _EMBEDDER(si.query)
            ctx = {"collection": str(collection_ids[0]), "vector_size": len(vector)}
            connector = VectorStoreConnectorAdaptor("b200", ctx).connector
            q = QueryWithEmbedding(query=si.query, top_k=ui.top_k, embedding=list(vector))
            results = connector.search(q, score_threshold=ui.similarity_threshold).results
            for item in results:
                if item.metadata is None:
                    item.metadata = {}
                item.metadata["recall_type"] = "vector_search"
            return VectorSearchOutput(docs=results), {}
        except Exception as e:  # runners/vector_search.py:103-105
            logging.getLogger(__name__).error(f"Vector search failed for collection {collection_ids[0]}: {e}")
            return VectorSearchOutput(docs=[]), {}


def install_connector() -> bool:
    """Make ``VECTOR_DB_TYPE=b200`` resolve inside an unmodified super-rag process: swap the
    adaptor class (super_rag/vectorstore/connector.py:4-15) for ours, which adds ``case "b200"`` and
    delegates every other type to the original. Returns False when the reference is not importable."""
    try:
        import sys
        import super_rag.vectorstore.connector as ref_mod  # type: ignore
    except Exception:
        return False
    from . import vectorstore
    if getattr(ref_mod.VectorStoreConnectorAdaptor, "_srag_b200", False):
        return True
    vectorstore._REFERENCE_ADAPTOR = ref_mod.VectorStoreConnectorAdaptor
    ref_mod.VectorStoreConnectorAdaptor = vectorstore.VectorStoreConnectorAdaptor
    ctx_mod = sys.modules.get("super_rag.context.context")
    if ctx_mod is not None:  # `from ... import VectorStoreConnectorAdaptor` bound the old class (context.py:4)
        ctx_mod.VectorStoreConnectorAdaptor = vectorstore.VectorStoreConnectorAdaptor
    _enable_filter_builder()
    return True


def _enable_filter_builder() -> None:
    """ContextManager builds its metadata filter only ``if self.vectordb_type == "seekdb"`` (context/context.py:60,
    92) and hands ``None`` to every other store. For type ``b200`` run the reference's OWN builder (unchanged) as if
    the store were seekdb, so the connector receives the filter document it can honour (ctx["apply_filter"])."""
    try:
        from super_rag.context.context import ContextManager  # type: ignore
    except Exception:
        return
    if getattr(ContextManager._create_combined_filter, "_srag_b200", False):
        return
    original = ContextManager._create_combined_filter

    def _create_combined_filter(self, index_types=None, chat_id=None):
        if getattr(self, "vectordb_type", None) != "b200":
            return original(self, index_types, chat_id)
        self.vectordb_type = "seekdb"
        try:
            return original(self, index_types, chat_id)
        finally:
            self.vectordb_type = "b200"

    _create_combined_filter._srag_b200 = True
    _create_combined_filter._original = original
    ContextManager._create_combined_filter = _create_combined_filter


def register() -> None:
    """Entry point of group ``super_rag.nodeflow.packs``."""
    register_node_runner("fulltext_search", input_model=FulltextSearchInput,
                         output_model=FulltextSearchOutput)(FulltextSearchNodeRunner)
    register_node_runner("merge", input_model=MergeInput, output_model=MergeOutput)(MergeNodeRunner)
    if "vector_search" not in NODE_RUNNER_REGISTRY:
        register_node_runner("vector_search", input_model=VectorSearchInput,
                             output_model=VectorSearchOutput)(B200VectorSearchNodeRunner)
    install_connector()

"""Data contract of the hot path. When the reference package is importable its own pydantic
models are used unchanged (super_rag/models/models.py:222-266,312-321); otherwise field-for-field
mirrors are defined here so the connector and nodes behave the same stand-alone."""
from __future__ import annotations

from typing import Any, Dict, List, Optional

try:  # pragma: no cover - exercised when running inside a super-rag checkout
    from super_rag.models import DocumentWithScore, Query, QueryResult, QueryWithEmbedding, TextNode  # type: ignore
    USING_REFERENCE_MODELS = True
except Exception:  # reference not on sys.path
    from pydantic import BaseModel

    USING_REFERENCE_MODELS = False

    class DocumentWithScore(BaseModel):  # models.py:263-266 (no id field: extra kwargs are dropped)
        text: Optional[str] = None
        score: Optional[float] = None
        metadata: Optional[dict] = None

    class Query(BaseModel):  # models.py:312-314
        query: str
        top_k: Optional[int] = 3

    class QueryWithEmbedding(Query):  # models.py:316-317
        embedding: List[float]

    class QueryResult(BaseModel):  # models.py:319-321
        query: str
        results: List[DocumentWithScore]

    class TextNode:  # models.py:222-256
        def __init__(self, text: str, metadata: Optional[Dict[str, Any]] = None, embedding: Optional[list] = None):
            self.text = text
            self.metadata = metadata or {}
            self.embedding = embedding

        def to_dict(self) -> Dict[str, Any]:
            return {"text": self.text, "metadata": self.metadata, "embedding": self.embedding}

        @classmethod
        def from_dict(cls, data: Dict[str, Any]) -> "TextNode":
            return cls(text=data.get("text", ""), metadata=data.get("metadata", {}), embedding=data.get("embedding"))

__all__ = ["DocumentWithScore", "Query", "QueryWithEmbedding", "QueryResult", "TextNode", "USING_REFERENCE_MODELS"]

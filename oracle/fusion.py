"""Oracle: candidate-list fusion (RRF / weighted / union merge / rerank fallback).

TEST INFRASTRUCTURE — see ``oracle/__init__.py`` for who may import this.

PINNED: ``rrf``, ``union_merge`` and ``rerank_fallback`` are checked against outputs of
the reference's own functions executed in this container (tests/golden/make_golden.py):
  * ``rrf``            <- super_rag/graphiti/graphiti_core/search/search_utils.py:1763-1778
  * ``union_merge``    <- super_rag/nodeflow/runners/merge.py:40-65
  * ``rerank_fallback``<- super_rag/nodeflow/runners/rerank.py:173-202
``rrf_fuse_topk`` / ``weighted_fuse_topk`` / ``merge_shard_candidates`` are the device
kernels' contract: same scores as ``rrf`` (float64, list-order accumulation) but the
north_star's order (score descending, chunk id ascending) and an optional top-k cut.
"""
from __future__ import annotations

from collections import defaultdict

import numpy as np


def rrf(results, rank_const=1, min_score: float = 0):
    """search_utils.py:1763-1778 restated: score[id] += 1/(i + rank_const), i 0-based;
    stable sort by score descending (ties keep first-seen order); keep score >= min_score."""
    scores = defaultdict(float)
    for result in results:
        for i, uid in enumerate(result):
            scores[uid] += 1 / (i + rank_const)
    scored = list(scores.items())
    scored.sort(reverse=True, key=lambda term: term[1])
    ids = [u for u, s in scored if s >= min_score]
    vals = [s for u, s in scored if s >= min_score]
    return ids, vals


def rrf_fuse_topk(lists_ids, rank_const: int = 1, min_score: float = 0.0, top_k: int | None = None):
    """Device contract for RRF. ``lists_ids``: sequence of 1-D integer id arrays (id < 0 =
    padding, skipped). Scores are float64 and identical to ``rrf``; order is
    (score desc, id asc). Returns (ids int64 [m], scores float64 [m]) with m <= top_k."""
    scores = {}
    for lst in lists_ids:
        for i, uid in enumerate(np.asarray(lst).tolist()):
            if uid < 0:
                continue
            scores[uid] = scores.get(uid, 0.0) + 1 / (i + rank_const)
    items = [(u, s) for u, s in scores.items() if s >= min_score]
    items.sort(key=lambda t: (-t[1], t[0]))
    if top_k is not None:
        items = items[:top_k]
    return (np.array([u for u, _ in items], dtype=np.int64),
            np.array([s for _, s in items], dtype=np.float64))


def weighted_fuse_topk(lists_ids, lists_scores, weights, top_k: int | None = None):
    """Weighted score fusion, fp32: acc = 0; for each occurrence in (list, position) order
    acc = fl(acc + fl(w_l * s)). Order (score desc, id asc)."""
    acc = {}
    for lst, sc, w in zip(lists_ids, lists_scores, weights):
        w = np.float32(w)
        for uid, s in zip(np.asarray(lst).tolist(), np.asarray(sc, dtype=np.float32)):
            if uid < 0:
                continue
            term = np.float32(w * np.float32(s))
            acc[uid] = np.float32(acc.get(uid, np.float32(0.0)) + term)
    items = sorted(acc.items(), key=lambda t: (-float(t[1]), t[0]))
    if top_k is not None:
        items = items[:top_k]
    return (np.array([u for u, _ in items], dtype=np.int64),
            np.array([s for _, s in items], dtype=np.float32))


def union_merge(doc_lists, deduplicate: bool = True, merge_strategy: str = "union"):
    """merge.py:40-65 restated on plain objects exposing ``.text`` (or dicts with "text").

    ``doc_lists`` is (vector, fulltext, graph, summary, vision) in that order (merge.py:56).
    Unknown strategy raises ValueError (the reference raises its own ValidationError)."""
    if merge_strategy not in ["union"]:
        raise ValueError(f"Unknown merge strategy: {merge_strategy}")
    all_docs = []
    for lst in doc_lists:
        all_docs = all_docs + list(lst or [])
    if not deduplicate:
        return all_docs
    seen = set()
    unique = []
    for doc in all_docs:
        text = doc["text"] if isinstance(doc, dict) else doc.text
        if text not in seen:
            seen.add(text)
            unique.append(doc)
    return unique


def rerank_fallback(docs):
    """rerank.py:173-202 restated: graph_search recalls first (input order), the rest
    stable-sorted by score descending with None -> 0.0."""
    if not docs:
        return docs

    def _get(doc, name):
        return doc[name] if isinstance(doc, dict) else getattr(doc, name)

    graph, other = [], []
    for doc in docs:
        meta = _get(doc, "metadata")
        if meta.get("recall_type", "") == "graph_search":
            graph.append(doc)
        else:
            other.append(doc)
    other.sort(key=lambda x: _get(x, "score") if _get(x, "score") is not None else 0.0, reverse=True)
    return graph + other


def merge_shard_candidates(scores: np.ndarray, ids: np.ndarray, k: int):
    """Final merge after the all-gather: scores fp32 [g,B,kk], ids int64 [g,B,kk]
    -> global top-k per query under (score desc, id asc); id < 0 entries are padding."""
    scores = np.asarray(scores, dtype=np.float32)
    ids = np.asarray(ids, dtype=np.int64)
    g, B, kk = scores.shape
    out_s = np.full((B, k), -np.inf, dtype=np.float32)
    out_i = np.full((B, k), -1, dtype=np.int64)
    for b in range(B):
        s = scores[:, b, :].reshape(-1)
        i = ids[:, b, :].reshape(-1)
        keep = i >= 0
        s, i = s[keep], i[keep]
        o = np.lexsort((i, -s.astype(np.float64)))[:k]
        out_s[b, :o.size] = s[o]
        out_i[b, :o.size] = i[o]
    return out_s, out_i

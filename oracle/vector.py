"""Oracle: exact cosine / inner-product top-k over chunk embeddings (numpy, CPU).

TEST INFRASTRUCTURE — see ``oracle/__init__.py`` for who may import this.

What it restates (reference paths relative to /root/reference):
  * ``SeekDBVectorStoreConnector.search`` -> ``collection.query(query_embeddings,
    n_results)`` under ``distance="cosine"``
    (super_rag/vectorstore/seekdb_connector.py:38,58-61,98-107). The arithmetic lives
    in the external SeekDB server (pyseekdb==1.0.0b8, pylibseekdb==1.0.1.post1,
    uv.lock:5747-5762,5550-5562; HNSW, approximate) and cannot run here. The oracle is the
    *exact* brute-force statement of that query and is PINNED to the reference's own in-tree
    brute-force cosine ranking — the Neptune branch of ``edge_similarity_search`` scores every
    stored embedding with ``calculate_cosine_similarity`` and orders by score descending
    (graphiti_core/search/search_utils.py:337-366 with :56-67): full ranked lists produced by
    that code, executed by ``tests/golden/make_golden.py``, are committed as
    ``tests/golden/topk_golden.json`` and asserted in ``tests/test_oracle_golden.py``
    (every pair score to 1e-14, ranked top-20 with duplicates, zero rows, a zero query,
    near-ties, the ``score > min_score`` cut).
  * zero-norm rule and the cosine formula follow ``calculate_cosine_similarity``
    (super_rag/graphiti/graphiti_core/search/search_utils.py:56-67):
    ``dot / (|a| * |b|)``, and 0 when either norm is 0.
  * ``normalize_l2`` follows super_rag/graphiti/graphiti_core/helpers.py:100-103
    (zero vector returned unchanged).
Ordering rule (north_star): score descending, chunk id ascending on ties.
"""
from __future__ import annotations

import numpy as np

METRIC_COSINE = 0
METRIC_IP = 1


def _metric_code(metric) -> int:
    if metric in (METRIC_COSINE, "cosine"):
        return METRIC_COSINE
    if metric in (METRIC_IP, "ip", "inner_product", "dot"):
        return METRIC_IP
    raise ValueError(f"unknown metric {metric!r}")


def bf16_round(x: np.ndarray) -> np.ndarray:
    """Round fp32 values to the nearest bf16 (ties to even); result is fp32 again.

    This is what the device index stores, so the oracle and the GPU consume the
    *same* inputs (SURVEY §7 "bit-exact ids with bf16 inputs").
    """
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    nan = np.isnan(x)
    lsb = (u >> 16) & 1
    r = ((u + 0x7FFF + lsb) >> 16) << 16
    out = (r & 0xFFFFFFFF).astype(np.uint32).view(np.float32).reshape(x.shape)
    if nan.any():
        out = out.copy()
        out[nan] = np.nan
    return out


def cosine_similarity_pair(a, b) -> float:
    """search_utils.py:56-67 restated (numpy float64 on python lists)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    dot = float(np.dot(a, b))
    na = float(np.linalg.norm(a))
    nb = float(np.linalg.norm(b))
    if na == 0 or nb == 0:
        return 0
    return dot / (na * nb)


def normalize_l2(v) -> np.ndarray:
    """helpers.py:100-103 restated."""
    arr = np.array(v)
    norm = np.linalg.norm(arr, 2, axis=0, keepdims=True)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(norm == 0, arr, arr / norm)


def row_scales(x: np.ndarray, metric="cosine") -> np.ndarray:
    """Per-row multiplier applied to the raw dot product: 1/|row| (0 for a zero row)
    for cosine, 1 for inner product. fp32, norm accumulated in fp64 then rounded."""
    x = np.asarray(x, dtype=np.float32)
    if _metric_code(metric) == METRIC_IP:
        return np.ones(x.shape[0], dtype=np.float32)
    n = np.empty(x.shape[0], dtype=np.float64)
    for r0 in range(0, x.shape[0], 65536):                     # bounded fp64 temporaries
        x64 = x[r0:r0 + 65536].astype(np.float64)
        n[r0:r0 + 65536] = np.sqrt(np.einsum("ij,ij->i", x64, x64))
    out = np.zeros(x.shape[0], dtype=np.float64)
    nz = n > 0
    out[nz] = 1.0 / n[nz]
    return out.astype(np.float32)


def exact_topk_blocks(blocks, queries: np.ndarray, k: int, metric="cosine", id_base: int = 0,
                      valid: np.ndarray | None = None):
    """Exact top-k of every query against a corpus delivered as an iterable of ``(first_row, fp32 block [m, D])``
    pairs (any order, disjoint): lets a 20 GB device corpus be checked at full size by streaming it to the host block
    by block. Same arithmetic as ``exact_topk``. ``valid`` is indexed by row."""
    queries = np.asarray(queries, dtype=np.float32)
    B = queries.shape[0]
    mc = _metric_code(metric)
    qs = row_scales(queries, mc)
    # running candidates as flat arrays (query, score, id); reduced to k per query after each block
    cq = np.zeros(0, dtype=np.int64)
    cs_ = np.zeros(0, dtype=np.float32)
    ci = np.zeros(0, dtype=np.int64)
    for r0, blk in blocks:
        blk = np.asarray(blk, dtype=np.float32)
        if blk.shape[0] == 0:
            continue
        dots = queries @ blk.T  # fp32 BLAS
        if mc == METRIC_COSINE:
            cs = row_scales(blk, mc)
            s = (dots * cs[None, :]) * qs[:, None]
        else:
            s = dots
        s = np.ascontiguousarray(s, dtype=np.float32)
        if valid is not None:
            s[:, ~valid[r0:r0 + blk.shape[0]]] = -np.inf
        n = s.shape[1]
        kk = min(k, n)
        kth = np.partition(s, n - kk, axis=1)[:, n - kk]      # k-th largest of each row
        rows, cols = np.nonzero(s >= kth[:, None])             # superset incl. ties at the k-th
        sc = s[rows, cols]
        keep = np.isfinite(sc)
        cq = np.concatenate([cq, rows[keep]])
        cs_ = np.concatenate([cs_, sc[keep]])
        ci = np.concatenate([ci, cols[keep].astype(np.int64) + r0 + id_base])
        order = np.lexsort((ci, -cs_.astype(np.float64), cq))   # query, score desc, id asc
        cq, cs_, ci = cq[order], cs_[order], ci[order]
        start = np.searchsorted(cq, np.arange(B), side="left")
        rank = np.arange(cq.shape[0]) - start[cq]
        top = rank < k
        cq, cs_, ci = cq[top], cs_[top], ci[top]
    out_s = np.full((B, k), -np.inf, dtype=np.float32)
    out_i = np.full((B, k), -1, dtype=np.int64)
    if cq.shape[0]:
        start = np.searchsorted(cq, np.arange(B), side="left")
        rank = np.arange(cq.shape[0]) - start[cq]
        out_s[cq, rank] = cs_
        out_i[cq, rank] = ci
    return out_s, out_i


def exact_topk(corpus: np.ndarray, queries: np.ndarray, k: int, metric="cosine",
               block_rows: int = 262144, id_base: int = 0, valid: np.ndarray | None = None):
    """Exact top-k of every query against every corpus row, fp32.

    score = (q . c) / (|q| |c|) for cosine (0 if either norm is 0), q . c for ip.
    Returns (scores fp32 [B,k], ids int64 [B,k]); rows short of k (N < k) are padded
    with (-inf, -1). ``valid`` (bool [N]) drops tombstoned rows.
    """
    corpus = np.asarray(corpus, dtype=np.float32)
    blocks = ((r0, corpus[r0:r0 + block_rows]) for r0 in range(0, corpus.shape[0], block_rows))
    return exact_topk_blocks(blocks, queries, k, metric, id_base, valid)


def exact_topk_scores_f64(corpus: np.ndarray, queries: np.ndarray, metric="cosine") -> np.ndarray:
    """All-pairs scores in float64 (the 'fp64 check variant' of SURVEY §8c) — small cases."""
    c = np.asarray(corpus, dtype=np.float64)
    q = np.asarray(queries, dtype=np.float64)
    s = q @ c.T
    if _metric_code(metric) == METRIC_COSINE:
        nq = np.linalg.norm(q, axis=1)
        nc = np.linalg.norm(c, axis=1)
        den = nq[:, None] * nc[None, :]
        with np.errstate(invalid="ignore", divide="ignore"):
            s = np.where(den == 0, 0.0, s / den)
    return s

"""Oracle: BM25 full-text scoring over an inverted index (numpy, CPU).

TEST INFRASTRUCTURE — see ``oracle/__init__.py`` for who may import this.

PARITY UNPINNED: the reference has no BM25 arithmetic in-tree. ``fulltext_search`` exists
only as a port / literal / view-model (super_rag/nodeflow/runners/merge.py:18-20,
super_rag/schema/view_models.py:1043-1047,1073-1081) and the design doc names "BM25"
(docs/design/nodeflow/nodeflow.md:83,94); vendored graphiti delegates to the graph DB's
Lucene index (graphiti_core/search/search_utils.py:562-574). This module restates the
published Okapi BM25 in the variant that KEEPS the ``(k1 + 1)`` numerator (Robertson's
formula, Lucene <= 7's BM25Similarity); Lucene >= 8 drops that constant factor, which leaves every
ranking unchanged and divides every score by k1 + 1 = 2.2:

    idf(t)   = ln(1 + (N - df + 0.5) / (df + 0.5))
    s(q, d)  = sum over unique t in q of  idf(t) * tf * (k1 + 1) / (tf + k1 * (1 - b + b * dl/avgdl))
    k1 = 1.2, b = 0.75

Numeric contract (shared with the device kernels so rankings are bit-exact):
  * idf is evaluated in float64 and rounded once to fp32;
  * every other step is ONE correctly-rounded fp32 operation, in exactly this order:
        r   = dl / avgdl          t = b * r        u = (1 - b) + t      nrm = k1 * u
        a   = tf * (k1 + 1)       num = idf * a    den = tf + nrm       imp = num / den
  * a document's score is the fp32 sum of its impacts added in ascending term-id order
    over the unique query terms, starting from +0.0f;
  * only documents matching at least one query term are candidates;
  * order: score descending, doc id ascending.
"""
from __future__ import annotations

import numpy as np

BM25_K1 = np.float32(1.2)
BM25_B = np.float32(0.75)
_ONE_MINUS_B = np.float32(1.0) - BM25_B          # exactly 0.25f
_K1_PLUS_1 = BM25_K1 + np.float32(1.0)           # fp32(1.2f + 1.0f)


def build_postings(tokens: np.ndarray, doc_ptr: np.ndarray, vocab: int):
    """Inverted index from flat token ids + CSR document boundaries.

    Returns term_ptr int64 [V+1], post_doc int32 [nnz] (ascending inside a term),
    post_tf int32 [nnz], dl int32 [N] (document lengths in tokens).
    """
    tokens = np.asarray(tokens, dtype=np.int64)
    doc_ptr = np.asarray(doc_ptr, dtype=np.int64)
    n_docs = doc_ptr.shape[0] - 1
    dl = np.diff(doc_ptr).astype(np.int32)
    doc_of_tok = np.repeat(np.arange(n_docs, dtype=np.int64), dl)
    key = tokens * n_docs + doc_of_tok
    uniq, counts = np.unique(key, return_counts=True)
    term = uniq // n_docs
    post_doc = (uniq % n_docs).astype(np.int32)
    post_tf = counts.astype(np.int32)
    df = np.bincount(term, minlength=vocab).astype(np.int64)
    term_ptr = np.zeros(vocab + 1, dtype=np.int64)
    np.cumsum(df, out=term_ptr[1:])
    return term_ptr, post_doc, post_tf, dl


def bm25_idf(df: np.ndarray, n_docs: int) -> np.ndarray:
    df = np.asarray(df, dtype=np.float64)
    return np.log(1.0 + (float(n_docs) - df + 0.5) / (df + 0.5)).astype(np.float32)


def bm25_avgdl(dl: np.ndarray) -> np.float32:
    """avgdl = float32(sum(dl) / N) with the division in float64."""
    dl = np.asarray(dl)
    return np.float32(float(dl.astype(np.int64).sum()) / float(max(1, dl.shape[0])))


def bm25_impacts(term_ptr, post_doc, post_tf, dl, n_docs_global=None, df_global=None,
                 avgdl=None) -> np.ndarray:
    """Per-posting impact, fp32, following the op order in the module docstring.

    ``n_docs_global`` / ``df_global`` / ``avgdl`` let a doc-range shard use the GLOBAL
    collection statistics so shard scores equal single-index scores bit for bit.
    """
    term_ptr = np.asarray(term_ptr, dtype=np.int64)
    V = term_ptr.shape[0] - 1
    df = np.diff(term_ptr) if df_global is None else np.asarray(df_global)
    n = int(dl.shape[0]) if n_docs_global is None else int(n_docs_global)
    idf = bm25_idf(df, n)
    avg = bm25_avgdl(dl) if avgdl is None else np.float32(avgdl)
    dlf = np.asarray(dl).astype(np.float32)
    r = dlf / avg
    t = BM25_B * r
    u = _ONE_MINUS_B + t
    nrm = (BM25_K1 * u).astype(np.float32)
    term_of_post = np.repeat(np.arange(V, dtype=np.int64), np.diff(term_ptr))
    tf = np.asarray(post_tf).astype(np.float32)
    a = tf * _K1_PLUS_1
    num = idf[term_of_post] * a
    den = tf + nrm[np.asarray(post_doc, dtype=np.int64)]
    return (num / den).astype(np.float32)


def bm25_scores_one(term_ptr, post_doc, post_imp, n_docs: int, query_terms):
    """Dense fp32 accumulator + matched mask for ONE query (ascending-term-id order)."""
    acc = np.zeros(n_docs, dtype=np.float32)
    hit = np.zeros(n_docs, dtype=bool)
    V = term_ptr.shape[0] - 1
    for t in sorted(set(int(x) for x in query_terms)):
        if t < 0 or t >= V:
            continue
        lo, hi = int(term_ptr[t]), int(term_ptr[t + 1])
        d = post_doc[lo:hi].astype(np.int64)
        acc[d] = acc[d] + post_imp[lo:hi]
        hit[d] = True
    return acc, hit


def bm25_topk(term_ptr, post_doc, post_imp, n_docs: int, queries, k: int, id_base: int = 0, valid=None):
    """Top-k per query. ``queries`` is a list of term-id sequences (negative ids = padding).
    ``valid``: optional bool [n_docs] — only these documents are candidates (metadata pre-filter; sums untouched).

    Returns (scores fp32 [B,k], ids int64 [B,k]) padded with (-inf, -1).
    """
    B = len(queries)
    out_s = np.full((B, k), -np.inf, dtype=np.float32)
    out_i = np.full((B, k), -1, dtype=np.int64)
    for b, q in enumerate(queries):
        acc, hit = bm25_scores_one(term_ptr, post_doc, post_imp, n_docs, q)
        if valid is not None:
            hit = hit & np.asarray(valid, dtype=bool)
        ids = np.nonzero(hit)[0].astype(np.int64)
        s = acc[ids]
        if ids.size > 4 * k + 64:
            kth = np.partition(s, ids.size - k)[ids.size - k]
            keep = s >= kth
            s, ids = s[keep], ids[keep]
        o = np.lexsort((ids, -s.astype(np.float64)))[:k]
        out_s[b, :o.size] = s[o]
        out_i[b, :o.size] = ids[o] + id_base
    return out_s, out_i

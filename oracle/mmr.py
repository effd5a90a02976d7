"""Maximal-marginal-relevance oracle. TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Restates ``maximal_marginal_relevance`` (graphiti_core/search/search_utils.py:1884-1922) in numpy float64.
PINNED: tests/test_oracle_golden.py checks it against outputs of the reference function itself
(tests/golden/mmr_golden.json, written by tests/golden/make_golden.py).

The reference is NOT the greedy textbook MMR; it is a one-shot score:

    c_i   = normalize_l2(candidate_i)                       (:1893-1895; a zero vector stays zero, helpers.py:100-103)
    S_ij  = c_i . c_j for i != j, S_ii = 0                   (:1899-1908; the matrix starts as zeros)
    mmr_i = lambda * (q . c_i) + (lambda - 1) * max_j S_ij   (:1911-1914; the max runs over the whole row, so it is
                                                              >= 0 because of the zero diagonal; q is not normalised)
    order = stable sort by mmr descending (:1916; ties keep candidate order), keep mmr_i >= min_score (:1921-1923)
"""
import numpy as np


def normalize_rows(C: np.ndarray) -> np.ndarray:
    C = np.asarray(C, dtype=np.float64)
    n = np.linalg.norm(C, axis=1, keepdims=True)
    return np.where(n == 0, C, C / np.where(n == 0, 1.0, n))


def mmr_scores(query: np.ndarray, candidates: np.ndarray, mmr_lambda: float = 0.5) -> np.ndarray:
    """float64 [m] scores in candidate order."""
    q = np.asarray(query, dtype=np.float64)
    C = normalize_rows(candidates)
    m = C.shape[0]
    if m == 0:
        return np.zeros(0)
    S = C @ C.T
    np.fill_diagonal(S, 0.0)
    return mmr_lambda * (C @ q) + (mmr_lambda - 1.0) * S.max(axis=1)


def mmr_rerank(query, candidates, mmr_lambda: float = 0.5, min_score: float = -2.0):
    """-> (order int64 [<= m] candidate positions, scores float64), sorted like the reference."""
    s = mmr_scores(query, candidates, mmr_lambda)
    order = np.argsort(-s, kind="stable")
    keep = s[order] >= min_score
    return order[keep].astype(np.int64), s[order][keep]

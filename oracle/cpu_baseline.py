"""Timed CPU baselines of the hot path (SURVEY §8d "CPU baseline timing"). TEST / BENCH INFRASTRUCTURE ONLY:
imported by ``bench.py`` (``cpu_baseline`` legs and ``--impl reference``) and nothing else.

  (ii)  vector: the best-effort CPU kernel — ``torch`` (MKL) fp32 ``Q @ C^T`` blocked over rows + ``torch.topk`` per
        block + a k-way merge of the running candidates, every host thread (``torch.set_num_threads`` set explicitly,
        because torchrun exports ``OMP_NUM_THREADS=1``). Same formula as ``oracle.exact_topk`` (cosine with precomputed
        inverse norms); the numpy oracle stays the CHECKER, this is the thing that is timed: its selection stage cost 5x
        the GEMM, ``torch.topk`` does not.
  (iii) BM25: numpy CSR gather + accumulate + ``argpartition`` per query (= ``oracle.bm25_topk``), one core.
  (iv)  fusion: the reference's own Python ``rrf`` restated (``oracle.rrf``, pinned to graphiti_core/search/
        search_utils.py:1763-1778) per query.
"""
from __future__ import annotations

import os
import time
from typing import Iterable, List, Sequence, Tuple


def cpu_model() -> str:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def use_all_threads() -> int:
    """Give BLAS every host thread this process may use and return the count actually set."""
    import torch
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def inv_norms(block):
    """1/|row| (0 for a zero row: search_utils.py:63-64), fp32."""
    import torch
    n = torch.linalg.vector_norm(block, dim=1)
    return torch.where(n > 0, 1.0 / n, torch.zeros_like(n))


def topk_step(blocks: Sequence, scales: Sequence, row0: Sequence[int], q, q_scale, k: int):
    """One pass of exact cosine top-k over fp32 row blocks (torch CPU tensors [m, D]) -> (scores [B,k], ids [B,k])."""
    import torch
    B = q.shape[0]
    best_s = torch.full((B, k), float("-inf"))
    best_i = torch.full((B, k), -1, dtype=torch.int64)
    for blk, sc, r0 in zip(blocks, scales, row0):
        s = q @ blk.T                                   # fp32 GEMM (MKL), all threads
        s *= sc[None, :]
        s *= q_scale[:, None]
        ts, ti = torch.topk(s, min(k, s.shape[1]), dim=1)
        cat_s = torch.cat([best_s, ts], dim=1)
        cat_i = torch.cat([best_i, ti + r0], dim=1)
        best_s, pos = torch.topk(cat_s, k, dim=1)       # k-way merge of the running candidates
        best_i = torch.gather(cat_i, 1, pos)
    return best_s, best_i


def time_topk(blocks, scales, row0, q, q_scale, k: int, steps: int, warmup: int) -> Tuple[float, tuple]:
    """Mean seconds per pass over the given blocks."""
    out = None
    for _ in range(warmup):
        out = topk_step(blocks, scales, row0, q, q_scale, k)
    t0 = time.perf_counter()
    for _ in range(steps):
        out = topk_step(blocks, scales, row0, q, q_scale, k)
    return (time.perf_counter() - t0) / max(1, steps), out


def time_topk_streamed(block_iter: Iterable, q, q_scale, k: int) -> Tuple[float, int, tuple]:
    """ONE un-extrapolated pass over a corpus delivered block by block (first_row, fp32 block): only the CPU compute is
    timed, fetching / generating the next block is not (the GPU arm's inputs are resident too).
    Returns (seconds of compute, rows scanned, (scores, ids))."""
    import torch
    B = q.shape[0]
    best_s = torch.full((B, k), float("-inf"))
    best_i = torch.full((B, k), -1, dtype=torch.int64)
    total, rows = 0.0, 0
    for r0, blk in block_iter:
        sc = inv_norms(blk)                             # ingest-time work (precomputed in the device index): not timed
        t0 = time.perf_counter()
        best_s, best_i = _merge_block(best_s, best_i, q, q_scale, blk, sc, r0, k)
        total += time.perf_counter() - t0
        rows += blk.shape[0]
    return total, rows, (best_s, best_i)


def _merge_block(best_s, best_i, q, q_scale, blk, sc, r0, k):
    import torch
    s = q @ blk.T
    s *= sc[None, :]
    s *= q_scale[:, None]
    ts, ti = torch.topk(s, min(k, s.shape[1]), dim=1)
    cat_s = torch.cat([best_s, ts], dim=1)
    cat_i = torch.cat([best_i, ti + r0], dim=1)
    best_s, pos = torch.topk(cat_s, k, dim=1)
    return best_s, torch.gather(cat_i, 1, pos)


def time_bm25(term_ptr, post_doc, post_imp, n_docs: int, queries: List[List[int]], k: int) -> float:
    """Mean seconds per query of the numpy CSR scorer (one core): gather the term's postings, accumulate, partition."""
    from . import bm25
    t0 = time.perf_counter()
    bm25.bm25_topk(term_ptr, post_doc, post_imp, n_docs, queries, k)
    return (time.perf_counter() - t0) / max(1, len(queries))


def time_rrf(list_pairs: List[Tuple[list, list]], rank_const: int, top_k: int) -> float:
    """Mean seconds per query of the reference's Python rrf (restated, pinned) + the top-k cut."""
    from . import fusion
    t0 = time.perf_counter()
    for a, b in list_pairs:
        ids, _ = fusion.rrf([a, b], rank_const=rank_const)
        ids[:top_k]
    return (time.perf_counter() - t0) / max(1, len(list_pairs))

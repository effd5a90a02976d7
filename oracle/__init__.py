"""CPU oracle for the hybrid-retrieval hot path. TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker or the
timed CPU baseline. Nothing under ``super_rag_b200/`` imports it; the product path
fails loudly when the CUDA library is missing.

Parity status:
  * fusion (rrf / union merge / rerank fallback / result wrapping): PINNED against
    outputs of the reference's own functions executed in this container
    (``tests/golden/make_golden.py`` -> ``tests/golden/*.json``).
  * cosine / inner-product top-k: PINNED against full ranked lists produced by the reference's
    own in-tree brute-force cosine ranking (``graphiti_core/search/search_utils.py:337-366``
    calling ``calculate_cosine_similarity``, ``:56-67``), executed in this container ->
    ``tests/golden/topk_golden.json``. (The production connector delegates the query to an
    external SeekDB server through ``pyseekdb==1.0.0b8``,
    ``super_rag/vectorstore/seekdb_connector.py:102-107``, absent from ``/root/reference`` and
    not installable here; the unmodified connector class is driven over a ``pyseekdb`` stub in
    ``tests/ref_stub.py``.)
  * BM25 arithmetic: PARITY UNPINNED and unpinnable — the reference has no BM25 arithmetic
    in-tree. The oracle restates Okapi BM25 with the ``(k1 + 1)`` numerator
    (Robertson / Lucene <= 7; Lucene >= 8 drops that constant factor: rankings identical, scores
    2.2x smaller).
"""
from .vector import (  # noqa: F401
    bf16_round,
    cosine_similarity_pair,
    normalize_l2,
    row_scales,
    exact_topk,
    exact_topk_blocks,
    exact_topk_scores_f64,
)
from .bm25 import (  # noqa: F401
    BM25_K1,
    BM25_B,
    bm25_idf,
    bm25_impacts,
    bm25_topk,
    build_postings,
)
from .fusion import (  # noqa: F401
    rrf,
    rrf_fuse_topk,
    weighted_fuse_topk,
    union_merge,
    rerank_fallback,
    merge_shard_candidates,
)
from .filters import eval_filter, filter_rows  # noqa: F401
from .mmr import mmr_rerank, mmr_scores  # noqa: F401

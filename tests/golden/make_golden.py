#!/usr/bin/env python
"""Generate golden vectors by EXECUTING the reference's own code in this container.

Run here (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

Nothing from the reference is copied into the repo: functions are loaded from the files
where they lie (``ast`` extraction for two modules whose imports need neo4j, path-imports
under stubbed parent modules for the nodeflow runners and the SeekDB connector), run on
seeded inputs, and only their INPUTS and OUTPUTS are written to ``tests/golden/*.json``.

Pinned here:
  rrf                        graphiti_core/search/search_utils.py:1763-1778
  calculate_cosine_similarity graphiti_core/search/search_utils.py:56-67
  normalize_l2               graphiti_core/helpers.py:100-103
  MergeNodeRunner.run        super_rag/nodeflow/runners/merge.py:40-65
  _apply_fallback_strategy   super_rag/nodeflow/runners/rerank.py:173-202
  SeekDBVectorStoreConnector.search / _convert_scored_point_to_document_with_score
                             super_rag/vectorstore/seekdb_connector.py:98-156
  maximal_marginal_relevance graphiti_core/search/search_utils.py:1884-1922  (-> mmr_golden.json)
  brute-force cosine RANKING  the Neptune branch of edge_similarity_search scores every stored embedding with
                             calculate_cosine_similarity and orders by score descending
                             (graphiti_core/search/search_utils.py:337-366 with :56-67)  (-> topk_golden.json)
"""
from __future__ import annotations

import ast
import asyncio
import importlib.util
import json
import os
import sys
import types
from collections import defaultdict

import numpy as np

REF = os.environ.get("SRAG_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def _extract_functions(path, names):
    src = open(path, encoding="utf-8").read()
    tree = ast.parse(src)
    out = {}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            out[node.name] = ast.get_source_segment(src, node)
    missing = set(names) - set(out)
    if missing:
        raise RuntimeError(f"{path}: missing {missing}")
    return out


def _load_graphiti_functions():
    from numpy.typing import NDArray
    ns = {"np": np, "defaultdict": defaultdict, "NDArray": NDArray}
    su = os.path.join(REF, "super_rag/graphiti/graphiti_core/search/search_utils.py")
    he = os.path.join(REF, "super_rag/graphiti/graphiti_core/helpers.py")
    for name, code in _extract_functions(su, ["rrf", "calculate_cosine_similarity"]).items():
        exec(compile(code, su, "exec"), ns)
    for name, code in _extract_functions(he, ["normalize_l2"]).items():
        exec(compile(code, he, "exec"), ns)
    return ns["rrf"], ns["calculate_cosine_similarity"], ns["normalize_l2"]


def _load_by_path(modname, relpath):
    spec = importlib.util.spec_from_file_location(modname, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


def _stub(modname, **attrs):
    mod = types.ModuleType(modname)
    mod.__dict__.update(attrs)
    mod.__path__ = []  # behave like a package for sub-imports
    sys.modules[modname] = mod
    return mod


def _load_runners():
    """Load the reference merge / rerank runners and the SeekDB connector unmodified."""
    sys.path.insert(0, REF)
    import super_rag.models as ref_models  # imports cleanly (pydantic only)

    registry = {}

    class BaseNodeRunner:  # stand-in for nodeflow/base/models.py:125-128 (needs langchain)
        pass

    class SystemInput:
        def __init__(self, query="", user="", **kw):
            self.query, self.user = query, user
            self.__dict__.update(kw)

    def register_node_runner(node_type, input_model, output_model):
        def deco(cls):
            registry[node_type] = {"runner": cls(), "input_model": input_model,
                                   "output_model": output_model}
            return cls
        return deco

    _stub("super_rag.nodeflow")
    _stub("super_rag.nodeflow.base")
    _stub("super_rag.nodeflow.base.models", BaseNodeRunner=BaseNodeRunner, SystemInput=SystemInput,
          register_node_runner=register_node_runner, NODE_RUNNER_REGISTRY=registry)
    _load_by_path("super_rag.nodeflow.base.exceptions", "super_rag/nodeflow/base/exceptions.py")
    _stub("super_rag.db")
    _stub("super_rag.db.ops", async_db_ops=None)
    _stub("super_rag.llm")

    class _E(Exception):
        pass

    _stub("super_rag.llm.llm_error_types", InvalidConfigurationError=_E, ProviderNotFoundError=_E,
          RerankError=_E)
    _stub("super_rag.llm.rerank")
    _stub("super_rag.llm.rerank.rerank_service", RerankService=object)
    merge = _load_by_path("ref_merge_runner", "super_rag/nodeflow/runners/merge.py")
    rerank = _load_by_path("ref_rerank_runner", "super_rag/nodeflow/runners/rerank.py")

    # pyseekdb stub (tests/ref_stub.py): only the three names imported at seekdb_connector.py:1-2,7; here in its
    # "canned" mode — every query returns the dict set through _Client.canned
    sys.path.insert(0, os.path.dirname(OUT))
    import ref_stub

    class _CannedSwitch:
        @property
        def canned(self):
            return ref_stub.STATE.canned

        @canned.setter
        def canned(self, value):
            ref_stub.STATE.canned = value

    _Client = _CannedSwitch()
    ref_stub.install()
    seek = _load_by_path("ref_seekdb_connector", "super_rag/vectorstore/seekdb_connector.py")
    return ref_models, merge, rerank, seek, _Client, SystemInput


def _doc_to_json(d):
    return {"text": d.text, "score": d.score, "metadata": d.metadata}


def _load_mmr():
    import logging
    from time import time
    from numpy.typing import NDArray
    su = os.path.join(REF, "super_rag/graphiti/graphiti_core/search/search_utils.py")
    he = os.path.join(REF, "super_rag/graphiti/graphiti_core/helpers.py")
    ns = {"np": np, "NDArray": NDArray, "time": time, "logger": logging.getLogger("ref"), "DEFAULT_MMR_LAMBDA": 0.5}
    for name, code in _extract_functions(he, ["normalize_l2"]).items():
        exec(compile(code, he, "exec"), ns)
    for name, code in _extract_functions(su, ["maximal_marginal_relevance"]).items():
        exec(compile(code, su, "exec"), ns)
    return ns["maximal_marginal_relevance"]


def make_mmr_golden():
    """Inputs are multiples of 1/64 in [-4, 4] (exact in bf16), so the GPU path sees the same numbers."""
    mmr = _load_mmr()
    rng = np.random.default_rng(77)
    cases = []
    for m, d, lam, ms in [(1, 8, 0.5, -2.0), (2, 8, 0.5, -2.0), (5, 16, 0.5, -2.0), (12, 32, 0.7, -2.0),
                          (33, 64, 0.3, 0.0), (64, 128, 0.5, 0.1), (100, 128, 0.5, -2.0), (200, 64, 0.9, -2.0),
                          (7, 16, 1.0, -2.0), (7, 16, 0.0, -2.0)]:
        q = (rng.integers(-256, 257, d) / 64.0).tolist()
        C = rng.integers(-256, 257, (m, d)) / 64.0
        if m >= 5:
            C[3] = C[1]                      # duplicate embedding: similarity 1, tie in the score
            C[4] = 0.0                       # zero vector: normalize_l2 leaves it unchanged
        cand = {f"u{i}": C[i].tolist() for i in range(m)}
        ids, sc = mmr(q, cand, lam, ms)
        cases.append({"query": q, "candidates": [C[i].tolist() for i in range(m)], "mmr_lambda": lam, "min_score": ms,
                      "order": [int(u[1:]) for u in ids], "scores": [float(x) for x in sc]})
    path = os.path.join(OUT, "mmr_golden.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump({"reference_commit": "49de377", "cases": cases}, f)
    print("wrote", path, len(cases))


def make_topk_golden():
    """Full ranked lists from the reference's own in-tree brute-force cosine ranking: the Neptune branch
    (search_utils.py:337-366) calls `calculate_cosine_similarity(search_vector, embedding)` (:56-67) for every row, keeps
    `score > min_score`, and orders by score descending with a limit. Here that loop runs over a seeded 2000 x 64 corpus
    with zero rows, exact duplicates, near-ties and a zero query; the ORDER BY is a stable sort, so equal scores stay in
    row order (= the north-star tie rule). Values are multiples of 1/64 in [-4, 4]: exact in bf16, so the GPU path
    sees the very numbers the reference function saw. Stored: the inputs (as integer numerators), every pair score
    (float64, as the function returned it) and the ranked top-20 per query."""
    _, cos, _ = _load_graphiti_functions()
    rng = np.random.default_rng(4242)
    n, d, k = 2000, 64, 20
    C = rng.integers(-256, 257, (n, d))
    C[17] = 0                                     # zero rows: score 0 whatever the query (search_utils.py:63-64)
    C[1999] = 0
    for a, b in [(5, 1200), (6, 7), (300, 1500), (1500, 1900)]:
        C[b] = C[a]                               # exact duplicates (a chain of three: 300, 1500, 1900)
    C[41] = C[40]
    C[41, 3] += 1                                 # near-ties: one numerator step in one coordinate
    C[900] = 2 * C[40]                            # same direction, twice the length: cosine equal up to rounding
    Q = rng.integers(-256, 257, (14, d))
    Q[0] = C[5]                                   # hits the duplicate pair 5 / 1200
    Q[1] = C[300]                                 # hits the triple
    Q[2] = C[40]                                  # near-tie neighbourhood
    Q[3] = 0                                      # zero query: every score is 0, ranking = row order
    Q[4] = -C[6]                                  # most scores negative around the duplicates 6 / 7
    Q[5] = C[123] + rng.integers(-8, 9, d)        # a planted neighbour
    min_scores = [-2.0] * 12 + [0.0, 0.2]         # the branch's `score > min_score` filter (last two queries)
    cases = []
    for qi in range(Q.shape[0]):
        qv = (Q[qi] / 64.0).tolist()
        scores = [float(cos(qv, (C[r] / 64.0).tolist())) for r in range(n)]
        kept = [{"id": r, "score": sc} for r, sc in enumerate(scores) if sc > min_scores[qi]]
        kept.sort(key=lambda e: -e["score"])      # ORDER BY i.score DESC (stable), LIMIT $limit
        top = kept[:k]
        cases.append({"query_num": Q[qi].tolist(), "min_score": min_scores[qi], "scores": scores,
                      "top_ids": [e["id"] for e in top], "top_scores": [e["score"] for e in top]})
    path = os.path.join(OUT, "topk_golden.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump({"reference_commit": "49de377", "denominator": 64, "k": k, "corpus_num": C.tolist(), "cases": cases}, f)
    print("wrote", path, len(cases))


def main():
    if "--mmr-only" in sys.argv:
        return make_mmr_golden()
    if "--topk-only" in sys.argv:
        return make_topk_golden()
    make_mmr_golden()
    make_topk_golden()
    rng = np.random.default_rng(20261104)
    rrf, cos, norm_l2 = _load_graphiti_functions()
    golden = {"reference_commit": "49de377", "cases": {}}

    # ---- rrf ----------------------------------------------------------------------
    rrf_cases = []
    fixed = [
        ([["a", "b", "c"], ["b", "a", "d"]], 1, 0),
        ([["a", "b", "c"], ["b", "a", "d"]], 1, 0.5),
        ([["a", "b", "c"], ["b", "a", "d"]], 60, 0),
        ([[], ["x"]], 1, 0),
        ([["x", "x", "y"], ["y"]], 1, 0),   # duplicate inside one list adds twice
        ([[]], 1, 0),
    ]
    for lists, rc, ms in fixed:
        ids, sc = rrf(lists, rank_const=rc, min_score=ms)
        rrf_cases.append({"lists": lists, "rank_const": rc, "min_score": ms, "ids": ids, "scores": sc})
    for trial in range(24):
        n_lists = int(rng.integers(1, 5))
        pool = int(rng.integers(5, 400))
        lists = []
        for _ in range(n_lists):
            ln = int(rng.integers(0, min(pool, 200) + 1))
            lists.append([int(x) for x in rng.choice(pool, size=ln, replace=False)])
        rc = int(rng.choice([1, 1, 2, 60]))
        ms = float(rng.choice([0.0, 0.0, 0.05, 0.5]))
        ids, sc = rrf(lists, rank_const=rc, min_score=ms)
        rrf_cases.append({"lists": lists, "rank_const": rc, "min_score": ms, "ids": ids, "scores": sc})
    golden["cases"]["rrf"] = rrf_cases

    # ---- cosine / normalize_l2 ------------------------------------------------------
    cos_cases = []
    for a, b in [([1, 0], [0, 0]), ([0, 0], [0, 0]), ([1, 0], [1, 0]), ([1, 0], [0, 1]),
                 ([1, 2, 3], [-1, -2, -3]), ([3, 4], [4, 3])]:
        cos_cases.append({"a": a, "b": b, "out": float(cos(a, b))})
    for _ in range(16):
        d = int(rng.integers(2, 64))
        a = [float(x) for x in rng.standard_normal(d)]
        b = [float(x) for x in rng.standard_normal(d)]
        cos_cases.append({"a": a, "b": b, "out": float(cos(a, b))})
    golden["cases"]["cosine"] = cos_cases
    nl_cases = []
    for v in [[0, 0], [3, 4], [0.0, 0.0, 0.0], [1e-3, -2e-3, 5.0]]:
        nl_cases.append({"v": v, "out": [float(x) for x in norm_l2(v)]})
    for _ in range(8):
        v = [float(x) for x in rng.standard_normal(int(rng.integers(1, 32)))]
        nl_cases.append({"v": v, "out": [float(x) for x in norm_l2(v)]})
    golden["cases"]["normalize_l2"] = nl_cases

    # ---- merge / rerank fallback / connector wrapping ------------------------------
    ref_models, merge, rerank, seek, Client, SystemInput = _load_runners()
    D = ref_models.DocumentWithScore
    runner = merge.MergeNodeRunner()
    si = SystemInput(query="q", user="u")

    def mk(prefix, n, recall, lo=0):
        return [{"text": f"{prefix}{i}", "score": float(np.round(rng.random(), 6)),
                 "metadata": {"recall_type": recall}} for i in range(lo, lo + n)]

    merge_cases = []
    specs = [
        dict(v=mk("t", 4, "vector_search"), f=mk("t", 4, "fulltext_search", lo=2), g=[], s=[], vi=[], dedup=True),
        dict(v=mk("t", 3, "vector_search"), f=mk("t", 3, "fulltext_search", lo=1), g=mk("g", 1, "graph_search"),
             s=mk("t", 2, "summary_search", lo=2), vi=mk("v", 2, "vision_search"), dedup=True),
        dict(v=mk("t", 3, "vector_search"), f=mk("t", 3, "fulltext_search", lo=1), g=[], s=[], vi=[], dedup=False),
        dict(v=[], f=[], g=[], s=[], vi=[], dedup=True),
        dict(v=mk("same", 1, "vector_search") + mk("same", 1, "vector_search"), f=[], g=[], s=[], vi=[], dedup=True),
    ]
    for sp in specs:
        ui = merge.MergeInput(merge_strategy="union", deduplicate=sp["dedup"],
                              vector_search_docs=[D(**d) for d in sp["v"]],
                              fulltext_search_docs=[D(**d) for d in sp["f"]],
                              graph_search_docs=[D(**d) for d in sp["g"]],
                              summary_search_docs=[D(**d) for d in sp["s"]],
                              vision_search_docs=[D(**d) for d in sp["vi"]])
        out, sysout = asyncio.run(runner.run(ui, si))
        merge_cases.append({"inputs": {k: sp[k] for k in ("v", "f", "g", "s", "vi")},
                            "deduplicate": sp["dedup"], "docs": [_doc_to_json(d) for d in out.docs],
                            "system_output": sysout})
    try:
        asyncio.run(runner.run(merge.MergeInput(merge_strategy="rrf"), si))
        unknown = None
    except Exception as e:  # reference: ValidationError("Unknown merge strategy: rrf")
        unknown = {"type": type(e).__name__, "message": str(e)}
    golden["cases"]["merge_union"] = merge_cases
    golden["cases"]["merge_unknown_strategy"] = unknown

    rr = rerank.RerankNodeRunner()
    fb_cases = []
    for trial in range(6):
        docs = mk("a", int(rng.integers(0, 6)), "vector_search") + mk("g", int(rng.integers(0, 3)), "graph_search") \
            + mk("b", int(rng.integers(0, 6)), "fulltext_search")
        if docs and trial % 2 == 0:
            docs[int(rng.integers(0, len(docs)))]["score"] = None
        if len(docs) >= 3:
            docs[1]["score"] = docs[0]["score"]  # tie: stable order must hold
        order = rng.permutation(len(docs))
        docs = [docs[i] for i in order]
        out = rr._apply_fallback_strategy([D(**d) for d in docs])
        fb_cases.append({"docs": docs, "out": [_doc_to_json(d) for d in out]})
    golden["cases"]["rerank_fallback"] = fb_cases

    conv_cases = []
    canned_list = [
        {"ids": [["i0", "i1"]], "distances": [[0.9, 0.5]], "documents": [["d0", "d1"]],
         "metadatas": [[{"a": 1}, {"b": 2}]]},
        {"ids": [["i0"]], "distances": [[None]], "documents": [["d0"]], "metadatas": None},
        {"ids": [["i0", "i1"], ["j0"]], "distances": [[0.3, 0.2], [0.1]], "documents": [["d0", "d1"], ["e0"]],
         "metadatas": [[None, {"k": "v"}], [None]]},
        {"ids": [], "distances": [], "documents": [], "metadatas": []},       # shape check fails -> None
        {"ids": "bad", "distances": None, "documents": None, "metadatas": None},
    ]
    for canned in canned_list:
        Client.canned = canned
        conn = seek.SeekDBVectorStoreConnector({"collection": "c"})
        q = ref_models.QueryWithEmbedding(query="hello", top_k=2, embedding=[0.0, 1.0])
        try:
            res = conn.search(q, score_threshold=0.3, filter={"x": 1})
            out = {"query": res.query, "results": [_doc_to_json(d) for d in res.results]}
        except Exception as e:  # results=None fails QueryResult validation in the reference
            out = {"raises": type(e).__name__}
        conv_cases.append({"hits": canned, "out": out})
    golden["cases"]["connector_wrap"] = conv_cases

    path = os.path.join(OUT, "fusion_golden.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump(golden, f, indent=1, sort_keys=True)
    print("wrote", path, {k: (len(v) if isinstance(v, list) else 1) for k, v in golden["cases"].items()})


if __name__ == "__main__":
    main()

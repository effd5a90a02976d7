"""Drop-in surface on the GPU: connector add / search / delete, the nodeflow nodes and a small hybrid
pipeline (vector top-k + BM25 top-k -> rrf merge), each compared with the oracle."""
import asyncio

import numpy as np
import pytest

import oracle
from parity import TOL, check_topk_against_full_scores

pytestmark = pytest.mark.gpu


def _run(coro):
    return asyncio.run(coro)


def _nodes(C, texts):
    from super_rag_b200.models import TextNode
    return [TextNode(text=t, metadata={"row": r}, embedding=C[r].tolist()) for r, t in enumerate(texts)]


def test_connector_add_search_delete_matches_oracle():
    from super_rag_b200.models import QueryWithEmbedding
    from super_rag_b200.vectorstore import VectorStoreConnectorAdaptor
    rng = np.random.default_rng(21)
    n, d = 700, 128
    C = rng.standard_normal((n, d)).astype(np.float32)
    C[300] = C[7]
    texts = [f"chunk {r}" for r in range(n)]
    name = "t_gpu_connector"
    con = VectorStoreConnectorAdaptor("b200", {"collection": name, "vector_size": d}).connector
    con.delete_collection()
    con.create_collection(vector_size=d)
    ids = con.add(_nodes(C, texts)[:400]) + con.add(_nodes(C, texts)[400:])     # two appends: growth path
    assert len(ids) == n and len(set(ids)) == n
    Cr = oracle.bf16_round(C)
    # a FRESH connector instance sees the same collection (the reference builds one per query)
    con2 = VectorStoreConnectorAdaptor("b200", {"collection": name, "vector_size": d}).connector
    for qi in (7, 123, 699):
        q = QueryWithEmbedding(query="hello", top_k=5, embedding=C[qi].tolist())
        res = con2.search(q, score_threshold=0.2, filter={"or": []}, limit=5, with_vectors=True)   # kwargs of context.py:37-47
        assert res.query == "hello" and len(res.results) == 5
        rows = [r.metadata["row"] for r in res.results]
        full = oracle.exact_topk_scores_f64(Cr, oracle.bf16_round(C[qi:qi + 1]), "cosine")
        check_topk_against_full_scores(np.array([[r.score for r in res.results]]), np.array([rows]), full, 5, TOL)
        assert res.results[0].text == texts[rows[0]]
    assert [r.metadata["row"] for r in con2.search(QueryWithEmbedding(query="q", top_k=2, embedding=C[7].tolist())).results] == [7, 300]
    # delete = tombstone: row 7 disappears, its duplicate 300 takes over
    con2.delete(ids=[ids[7]])
    res = con2.search(QueryWithEmbedding(query="q", top_k=3, embedding=C[7].tolist()))
    assert res.results[0].metadata["row"] == 300 and 7 not in [r.metadata["row"] for r in res.results]
    assert con2.search(QueryWithEmbedding(query="q", top_k=0, embedding=C[0].tolist())).results == []
    con2.delete_collection()
    assert con.search(QueryWithEmbedding(query="q", top_k=3, embedding=C[0].tolist())).results == []


def test_fulltext_node_matches_bm25_oracle():
    from super_rag_b200 import fulltext, nodeflow_pack as n
    rng = np.random.default_rng(5)
    words = [f"w{i}" for i in range(400)]
    p = 1.0 / np.arange(1, 401)
    p /= p.sum()
    texts = [" ".join(rng.choice(words, size=int(rng.integers(5, 40)), p=p)) for _ in range(3000)]
    enc = fulltext.hash_tokenizer(4096)
    fulltext.delete_collection("t_gpu_ft")
    col = fulltext.get_or_create_collection("t_gpu_ft", tokenizer=enc, vocab=4096)
    col.add(texts, [{"row": r} for r in range(len(texts))])
    n.register()
    runner = n.NODE_RUNNER_REGISTRY["fulltext_search"]["runner"]
    # oracle index over the same tokens
    toks = [np.asarray(enc(t), dtype=np.int64) for t in texts]
    doc_ptr = np.concatenate([[0], np.cumsum([len(t) for t in toks])]).astype(np.int64)
    tp, pd, ptf, dl = oracle.build_postings(np.concatenate(toks), doc_ptr, 4096)
    imp = oracle.bm25_impacts(tp, pd, ptf, dl)
    for query, kw in [("w3 w17 w250", None), ("ignored", ["w1", "w2", "w399"]), ("nothing-matches-this", None)]:
        si = n.SystemInput(query=query, user="u", collection_ids=["t_gpu_ft"])
        out, _ = _run(runner.run(n.FulltextSearchInput(top_k=20, keywords=kw), si))
        terms = enc(" ".join(kw) if kw else query)
        os_, oi = oracle.bm25_topk(tp, pd, imp, len(texts), [terms], 20)
        want = [(int(i), float(s)) for s, i in zip(os_[0], oi[0]) if i >= 0]
        got = [(d.metadata["row"], d.score) for d in out.docs]
        assert got == want, (query, got[:3], want[:3])
        assert all(d.metadata["recall_type"] == "fulltext_search" and d.text == texts[d.metadata["row"]] for d in out.docs)
    fulltext.delete_collection("t_gpu_ft")


def test_merge_rrf_matches_reference_golden(golden):
    """merge_strategy='rrf' on the GPU reproduces the reference rrf() (ids, float64 scores, first-seen ties)."""
    from super_rag_b200 import nodeflow_pack as n
    from super_rag_b200.models import DocumentWithScore as D
    n.register()
    runner = n.NODE_RUNNER_REGISTRY["merge"]["runner"]
    names = ["vector_search_docs", "fulltext_search_docs", "graph_search_docs", "summary_search_docs",
             "vision_search_docs"]
    done = 0
    for c in golden["rrf"]:
        if len(c["lists"]) > 5 or any(len(set(l)) != len(l) for l in c["lists"]):
            continue        # documents are identified by text: a text repeated inside one list is one document
        kw = {nm: [D(text=str(x), score=0.5) for x in l] for nm, l in zip(names, c["lists"])}
        out, _ = _run(runner.run(n.MergeInput(merge_strategy="rrf", rank_const=c["rank_const"],
                                               min_score=c["min_score"], **kw), None))
        assert [d.text for d in out.docs] == [str(x) for x in c["ids"]]
        assert [d.score for d in out.docs] == c["scores"]
        done += 1
    assert done >= 10
    out, _ = _run(runner.run(n.MergeInput(merge_strategy="rrf", top_k=1, vector_search_docs=[D(text="a"), D(text="b")],
                                           fulltext_search_docs=[D(text="b")]), None))
    assert [d.text for d in out.docs] == ["b"] and out.docs[0].score == 1.0 / 2 + 1.0
    out, _ = _run(runner.run(n.MergeInput(merge_strategy="weighted", weights=[1.0, 2.0, 0, 0, 0],
                                           vector_search_docs=[D(text="a", score=0.5), D(text="b", score=0.25)],
                                           fulltext_search_docs=[D(text="b", score=0.5)]), None))
    assert [(d.text, d.score) for d in out.docs] == [("b", 1.25), ("a", 0.5)]


def test_hybrid_pipeline_small():
    """C4 in miniature: vector top-50 + BM25 top-50 -> RRF -> top-10, ids vs oracle, through the batched API."""
    import torch
    from super_rag_b200 import ops
    from super_rag_b200.bm25 import Bm25Index
    rng = np.random.default_rng(9)
    n, d, b, vocab = 20000, 256, 64, 3000
    C = rng.standard_normal((n, d)).astype(np.float32)
    Q = C[rng.integers(0, n, b)] + 0.3 * rng.standard_normal((b, d)).astype(np.float32)
    dev = torch.device("cuda:0")
    cb, cs = ops.ingest_rows(torch.from_numpy(C).to(dev), "cosine")
    qb, qs = ops.ingest_rows(torch.from_numpy(Q).to(dev), "cosine")
    vs, vi = ops.topk_search(cb, cs, qb, qs, 50)
    dl = rng.integers(30, 90, n)
    p = 1.0 / np.arange(1, vocab + 1)
    tokens = rng.choice(vocab, size=int(dl.sum()), p=p / p.sum()).astype(np.int64)
    doc_ptr = np.concatenate([[0], np.cumsum(dl)]).astype(np.int64)
    idx = Bm25Index.from_tokens(torch.from_numpy(tokens).to(dev), torch.from_numpy(doc_ptr).to(dev), vocab)
    queries = [rng.integers(50, vocab, 8).tolist() for _ in range(b)]
    bs, bi = idx.search(queries, 50)
    for rc in (1, 60):
        fi, fs = ops.fuse(torch.stack([vi, bi]), None, ops.FUSE_RRF, 10, rank_const=rc)
        torch.cuda.synchronize()
        for q in range(b):
            wi, ws = oracle.rrf_fuse_topk([vi[q].cpu().numpy(), bi[q].cpu().numpy()], rc, 0.0, 10)
            assert fi[q].cpu().numpy()[:len(wi)].tolist() == wi.tolist()
            assert fs[q].cpu().numpy()[:len(ws)].tolist() == ws.tolist()
    # and the two recalls themselves
    full = oracle.exact_topk_scores_f64(oracle.bf16_round(C), oracle.bf16_round(Q), "cosine")
    check_topk_against_full_scores(vs.cpu().numpy(), vi.cpu().numpy(), full, 50, TOL)
    tp, pd, ptf, dl32 = oracle.build_postings(tokens, doc_ptr, vocab)
    obs, obi = oracle.bm25_topk(tp, pd, oracle.bm25_impacts(tp, pd, ptf, dl32), n, queries, 50)
    assert (bi.cpu().numpy() == obi).all() and (bs.cpu().numpy().view(np.uint32) == obs.view(np.uint32)).all()

"""Metadata pre-filter folded into K1: the result must be the exact top-k of the rows the
reference's filter document admits (oracle.eval_filter on the raw metadata), through the connector API with
ctx["apply_filter"], the batched API with per-query filters, and the raw C-ABI entry point."""
import numpy as np
import pytest

import oracle
from parity import TOL, check_topk_against_full_scores

pytestmark = pytest.mark.gpu

INDEXERS = ["vector", "vision", "summary", None]        # None = legacy chunk without the field


def _metas(rng, n):
    out = []
    for r in range(n):
        m = {"row": r}
        ix = INDEXERS[int(rng.integers(0, 4))]
        if ix is not None:
            m["indexer"] = ix
        if rng.random() < 0.3:
            m["chat_id"] = f"chat-{int(rng.integers(0, 5))}"
        out.append(m)
    return out


def _filter(index_types=None, chat_id=None):
    """What ContextManager._create_combined_filter returns (context/context.py:74-111)."""
    clauses = []
    if index_types:
        clauses.append({"or": [{"indexer": {"$in": index_types}}, {"indexer": {"$exists": False}}]})
    if chat_id:
        clauses.append({"chat_id": chat_id})
    if not clauses:
        return None
    return clauses[0] if len(clauses) == 1 else {"and": clauses}


def _check(scores, rows, C_r, q_r, admit, k):
    full = oracle.exact_topk_scores_f64(C_r, q_r, "cosine")
    full[:, ~np.asarray(admit)] = -np.inf
    n_ok = int(np.sum(admit))
    kk = min(k, n_ok)
    rows = np.asarray(rows)
    assert (rows[:kk] >= 0).all() and (rows[kk:] < 0).all(), (rows, n_ok)
    assert all(admit[r] for r in rows[:kk])
    if kk:
        check_topk_against_full_scores(np.asarray(scores)[None, :kk], rows[None, :kk], full, kk, TOL)


def test_filtered_search_matches_oracle():
    import torch
    from super_rag_b200.models import QueryWithEmbedding, TextNode
    from super_rag_b200.vectorstore import VectorStoreConnectorAdaptor
    rng = np.random.default_rng(31)
    n, d, k = 6000, 128, 8
    C = rng.standard_normal((n, d)).astype(np.float32)
    metas = _metas(rng, n)
    name = "t_gpu_filter"
    ctx = {"collection": name, "vector_size": d, "apply_filter": True}
    con = VectorStoreConnectorAdaptor("b200", ctx).connector
    con.delete_collection()
    con.create_collection(vector_size=d)
    con.add([TextNode(text=f"c{r}", metadata=metas[r], embedding=C[r].tolist()) for r in range(n)])
    C_r = oracle.bf16_round(C)
    filters = [_filter(["vector"]), _filter(["vision", "summary"]), _filter(None, "chat-2"),
               _filter(["summary"], "chat-4"), _filter(["nonexistent"]), _filter(None, "no-such-chat"), None]
    # single-query connector API (the call ContextManager.query makes, context.py:37-47)
    for fi, flt in enumerate(filters):
        qi = 17 * fi + 3
        q = C[qi] + 0.05 * rng.standard_normal(d).astype(np.float32)
        res = con.search(QueryWithEmbedding(query="q", top_k=k, embedding=q.tolist()), filter=flt, limit=k)
        admit = oracle.filter_rows(flt, metas)
        rows = [r.metadata["row"] for r in res.results]
        assert len(rows) == min(k, sum(admit))
        _check([r.score for r in res.results], rows + [-1] * (k - len(rows)), C_r, oracle.bf16_round(q[None]), admit, k)
    # the reference behaviour (filter dropped) stays the default
    con_plain = VectorStoreConnectorAdaptor("b200", {"collection": name, "vector_size": d}).connector
    res = con_plain.search(QueryWithEmbedding(query="q", top_k=3, embedding=C[5].tolist()), filter=_filter(None, "no-such-chat"))
    assert [r.metadata["row"] for r in res.results][0] == 5
    # batched API: a different filter per query, B > 128 so several q-tiles and both CTA-pair plans are used
    B = 300
    Q = C[rng.integers(0, n, B)] + 0.05 * rng.standard_normal((B, d)).astype(np.float32)
    per_q = [filters[i % len(filters)] for i in range(B)]
    s, i = con.search_batch(torch.from_numpy(Q), k, filters=per_q)
    s, i = s.cpu().numpy(), i.cpu().numpy()
    Q_r = oracle.bf16_round(Q)
    for b in range(0, B, 7):
        _check(s[b], i[b], C_r, Q_r[b:b + 1], oracle.filter_rows(per_q[b], metas), k)
    # a tombstone and a filter together
    col = con._collection()
    first = next(r for r in range(n) if metas[r].get("chat_id") == "chat-2")
    col.delete([col.docs.ids[first]])
    res = con.search(QueryWithEmbedding(query="q", top_k=k, embedding=C[first].tolist()), filter=_filter(None, "chat-2"))
    assert first not in [r.metadata["row"] for r in res.results] and len(res.results) == k
    con.delete_collection()


def test_filtered_c_abi_small_k_and_groups_only():
    import torch
    from super_rag_b200 import ops
    rng = np.random.default_rng(2)
    n, d, b, k = 20000, 64, 40, 5
    dev = torch.device("cuda:0")
    C = rng.standard_normal((n, d)).astype(np.float32)
    Q = rng.standard_normal((b, d)).astype(np.float32)
    cb, cs = ops.ingest_rows(torch.from_numpy(C).to(dev), "cosine")
    qb, qs = ops.ingest_rows(torch.from_numpy(Q).to(dev), "cosine")
    group = rng.integers(0, 50, n).astype(np.int32)            # ~400 rows per group: a selective filter
    qg = rng.integers(-1, 50, b).astype(np.int32)
    s, i = ops.topk_search(cb, cs, qb, qs, k, row_group=torch.from_numpy(group).to(dev), q_group=torch.from_numpy(qg).to(dev))
    s, i = s.cpu().numpy(), i.cpu().numpy()
    C_r, Q_r = oracle.bf16_round(C), oracle.bf16_round(Q)
    for q in range(b):
        admit = np.ones(n, bool) if qg[q] < 0 else group == qg[q]
        _check(s[q], i[q], C_r, Q_r[q:q + 1], admit, k)
    with pytest.raises(Exception):
        ops.topk_search(cb, cs, qb, qs, k, row_group=torch.from_numpy(group).to(dev))


def test_bm25_filtered_matches_oracle_small_and_two_phase():
    """K2 with the metadata pre-filter: exact BM25 top-k among the admitted documents only (scores bit-exact:
    the filter removes candidates, it never changes a sum). Covers the single-phase path (few blocks) and the
    sampled two-phase path (>= 32 blocks) with a selective and a permissive filter."""
    import torch
    from super_rag_b200.bm25 import Bm25Index
    dev = torch.device("cuda:0")
    for n_docs, vocab, k in [(40_000, 500, 20), (600_000, 300, 50)]:
        rng = np.random.default_rng(n_docs)
        dl = rng.integers(5, 30, n_docs)
        p = 1.0 / np.arange(1, vocab + 1)
        tokens = rng.choice(vocab, size=int(dl.sum()), p=p / p.sum()).astype(np.int64)
        doc_ptr = np.concatenate([[0], np.cumsum(dl)]).astype(np.int64)
        idx = Bm25Index.from_tokens(torch.from_numpy(tokens).to(dev), torch.from_numpy(doc_ptr).to(dev), vocab)
        tp, pd, ptf, dl32 = oracle.build_postings(tokens, doc_ptr, vocab)
        imp = oracle.bm25_impacts(tp, pd, ptf, dl32)
        tags = (1 << rng.integers(0, 4, n_docs)).astype(np.int32)          # one of 4 categories per document
        group = rng.integers(0, 200, n_docs).astype(np.int32)             # 200 groups: a selective constraint
        queries = [rng.integers(0, vocab, 6).tolist() for _ in range(12)]
        q_mask = np.array([0b0011, 0b1000, 0b1111, 0b0100] * 3, dtype=np.int32)
        q_group = np.array([-1, -1, 7, 7, 150, -1] * 2, dtype=np.int32)
        s, i = idx.search(queries, k, doc_tags=torch.from_numpy(tags).to(dev), q_tag_mask=torch.from_numpy(q_mask).to(dev),
                          doc_group=torch.from_numpy(group).to(dev), q_group=torch.from_numpy(q_group).to(dev))
        s, i = s.cpu().numpy(), i.cpu().numpy()
        # oracle: every match of the query, restricted to the admitted documents, (score desc, id asc)
        from oracle.bm25 import bm25_scores_one
        for q in range(len(queries)):
            acc, hit = bm25_scores_one(tp, pd, imp, n_docs, queries[q])
            admit = hit & ((tags & q_mask[q]) != 0) & ((q_group[q] < 0) | (group == q_group[q]))
            ids_q = np.nonzero(admit)[0]
            order = np.lexsort((ids_q, -acc[ids_q].astype(np.float64)))[:k]
            want_i, want_s = ids_q[order], acc[ids_q][order].astype(np.float32)
            n = len(want_i)
            assert i[q, :n].tolist() == want_i.tolist(), (n_docs, q)
            assert (s[q, :n].view(np.uint32) == want_s.view(np.uint32)).all()
            assert (i[q, n:] == -1).all()


def test_fulltext_collection_chat_filter():
    from super_rag_b200 import fulltext
    enc = fulltext.hash_tokenizer(4096)
    col = fulltext.FulltextCollection("t_ft_filter", tokenizer=enc, vocab=4096, apply_filter=True)
    texts = [f"alpha common w{r % 30}" for r in range(2000)]
    metas = [{"row": r, **({"chat_id": f"c{r % 4}"} if r % 2 else {})} for r in range(2000)]
    col.add(texts, metas)
    plain = col.search("alpha w3", 10)
    only = col.search("alpha w3", 10, filter={"chat_id": "c1"})
    assert len(plain) == 10 and len(only) == 10
    assert all(d.metadata.get("chat_id") == "c1" for d in only)
    assert any(d.metadata.get("chat_id") != "c1" for d in plain)
    assert col.search("alpha w3", 10, filter={"chat_id": "nobody"}) == []
    ignoring = fulltext.FulltextCollection("t_ft_nofilter", tokenizer=enc, vocab=4096)       # default: filter ignored
    ignoring.add(texts, metas)
    assert [d.text for d in ignoring.search("alpha w3", 10, filter={"chat_id": "c1"})] == [d.text for d in plain]

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

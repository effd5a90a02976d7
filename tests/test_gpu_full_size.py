"""BASELINE.json's full sizes, checked through size-independent properties (the oracle would need hours):

  C2  10M x 1024 bf16, 1024 queries, top-10: planted neighbours come back first with cosine ~ 1,
      rows sorted by (score desc, id asc), no duplicates, scores reproduce when the returned rows are
      re-scored in fp64, decomposition (top-k of the whole == merge of the top-k of two halves, bit for
      bit) and idempotence.
  C3  BM25 over 10M documents: the same decomposition over document ranges with global statistics is
      covered at 120k docs over NCCL (tools/check_sharded_nccl.py); here the block-sampled two-phase
      selection is compared with an independent torch restatement on 2M docs x 50k vocab.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_c1_connector_full_size_vs_oracle():
    """C1: 100k x 1024 fp32 embeddings, top-5, through the connector API (single-query calls and the batch
    call), against the fp64 oracle on the bf16-rounded inputs the device stores."""
    import torch
    import oracle
    from parity import TOL, check_topk_against_full_scores
    from super_rag_b200.models import QueryWithEmbedding
    from super_rag_b200.vectorstore import B200Client, VectorStoreConnectorAdaptor
    n, d, nq, k = 100_000, 1024, 256, 5
    rng = np.random.default_rng(1234)
    C = rng.standard_normal((n, d)).astype(np.float32)
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    src = rng.integers(0, n, nq // 2)
    Q[nq // 2:] = C[src] + 0.1 * rng.standard_normal((nq // 2, d)).astype(np.float32)
    B200Client().delete_collection("t_c1")
    col = B200Client().create_collection("t_c1", d)
    col.append(torch.from_numpy(C), metas=[{"row": r} for r in range(n)])
    con = VectorStoreConnectorAdaptor("b200", {"collection": "t_c1", "vector_size": d}).connector
    s, i = con.search_batch(torch.from_numpy(Q), k)
    s, i = s.cpu().numpy(), i.cpu().numpy()
    full = oracle.exact_topk_scores_f64(oracle.bf16_round(C), oracle.bf16_round(Q), "cosine")
    check_topk_against_full_scores(s, i, full, k, TOL)
    assert (i[nq // 2:, 0] == src).all()
    for b in (0, 1, nq // 2, nq - 1):       # the single-query API returns the batch call's row
        res = con.search(QueryWithEmbedding(query="q", top_k=k, embedding=Q[b].tolist()))
        assert [r.metadata["row"] for r in res.results] == i[b].tolist()
        assert np.allclose([r.score for r in res.results], s[b], atol=1e-6)
    B200Client().delete_collection("t_c1")


def _corpus(torch, n, d, dev):
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    out = torch.empty(n, d, dtype=torch.bfloat16, device=dev)
    for r0 in range(0, n, 1 << 20):
        r1 = min(n, r0 + (1 << 20))
        out[r0:r1] = torch.randn(r1 - r0, d, generator=g, device=dev).to(torch.bfloat16)
    return out, g


def test_c2_full_size_properties():
    import torch
    from super_rag_b200 import ops
    dev = torch.device("cuda:0")
    if torch.cuda.get_device_properties(dev).total_memory < 60e9:
        pytest.skip("needs a GPU with > 60 GB")
    n, d, b, k = 10_000_000, 1024, 1024, 10
    corpus, g = _corpus(torch, n, d, dev)
    _, cs = ops.ingest_rows(corpus, "cosine")
    q = torch.randn(b, d, generator=g, device=dev)
    h = b // 2
    src = torch.randint(0, n, (h,), generator=g, device=dev)
    q[:h] = corpus[src].float() + 0.05 * q[:h]
    # exact duplicates far apart: the lower id must win the tie
    corpus[9_000_000:9_000_064] = corpus[1000:1064]
    cs[9_000_000:9_000_064] = cs[1000:1064]
    q[h:h + 64] = corpus[1000:1064].float()
    qb, qs = ops.ingest_rows(q, "cosine")
    s, i = ops.topk_search(corpus, cs, qb, qs, k)
    torch.cuda.synchronize()
    # planted neighbours first, duplicates ordered by id
    assert torch.equal(i[:h, 0], src)
    assert (s[:h, 0] > 0.99).all()
    assert torch.equal(i[h:h + 64, 0], torch.arange(1000, 1064, device=dev))
    assert torch.equal(i[h:h + 64, 1], torch.arange(9_000_000, 9_000_064, device=dev))
    assert torch.equal(s[h:h + 64, 0].view(torch.int32), s[h:h + 64, 1].view(torch.int32))
    # order, uniqueness, range
    assert (i >= 0).all() and (i < n).all()
    assert (s[:, :-1] >= s[:, 1:]).all()
    tie = s[:, :-1] == s[:, 1:]
    assert (i[:, :-1][tie] < i[:, 1:][tie]).all()
    assert (torch.sort(i, dim=1).values[:, 1:] != torch.sort(i, dim=1).values[:, :-1]).all()
    # returned scores reproduce in fp64 from the bf16 inputs (1e-3 is the north-star tolerance)
    rows = corpus[i.reshape(-1)].double().view(b, k, d)
    qq = qb.double()
    ref = (rows * qq[:, None, :]).sum(-1) / (rows.norm(dim=-1) * qq.norm(dim=-1)[:, None])
    assert (ref - s.double()).abs().max().item() < 1e-3
    # decomposition: whole == merge(top-k of halves); idempotence
    mid = 5_000_000
    s1, i1 = ops.topk_search(corpus[:mid], cs[:mid], qb, qs, k)
    s2, i2 = ops.topk_search(corpus[mid:], cs[mid:], qb, qs, k, id_base=mid)
    ms, mi = ops.merge_candidates(torch.stack([s1, s2]), torch.stack([i1, i2]), k)
    assert torch.equal(mi, i) and torch.equal(ms.view(torch.int32), s.view(torch.int32))
    s3, i3 = ops.topk_search(corpus, cs, qb, qs, k)
    assert torch.equal(i3, i) and torch.equal(s3.view(torch.int32), s.view(torch.int32))
    # k-th score really is a frontier: no row of a random 200k-row sample beats it for 8 queries
    sample = torch.randint(0, n, (200_000,), generator=g, device=dev)
    sc = (corpus[sample].float() @ qb[:8].float().T) * cs[sample][:, None] * qs[:8][None, :]
    assert (sc.max(dim=0).values <= s[:8, 0] + 1e-3).all()
    better = sc > (s[:8, k - 1][None, :] + 1e-3)
    for j in range(8):
        assert set(sample[better[:, j]].tolist()) <= set(i[j].tolist())


def test_c3_two_phase_selection_matches_restatement():
    import os
    import sys
    import torch
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import bench_hybrid as bh
    dev = torch.device("cuda:0")
    idx = bh.build_index(2_000_000, dev)            # 123 blocks: phase 0 samples every 32nd, phase 1 runs on the warp-granular kernel
    qt = bh.make_queries(48, 8, dev)
    qt[0, 4:] = -1                                   # short query
    qt[1] = torch.tensor([0, 1, 2, 3, 4, 5, 6, 7], dtype=torch.int32, device=dev)   # the 8 most frequent terms
    s, i = idx.search_terms(qt, 100)
    torch.cuda.synchronize()
    for q in range(0, 48, 3):
        ws, wi = bh.torch_restatement(idx, qt[q].tolist(), 100)
        assert torch.equal(i[q], wi) and torch.equal(s[q].view(torch.int32), ws.view(torch.int32)), q

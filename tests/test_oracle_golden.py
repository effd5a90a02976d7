"""Pin the oracle against outputs of the reference's own code (tests/golden/make_golden.py)."""
import numpy as np
import pytest

import oracle


class _Doc:
    def __init__(self, d):
        self.text, self.score, self.metadata = d["text"], d["score"], d["metadata"]

    def to_json(self):
        return {"text": self.text, "score": self.score, "metadata": self.metadata}


def test_rrf_matches_reference(golden):
    for c in golden["rrf"]:
        ids, sc = oracle.rrf(c["lists"], rank_const=c["rank_const"], min_score=c["min_score"])
        assert ids == c["ids"]
        assert sc == c["scores"]  # float64, bit-exact


def test_rrf_device_contract_same_scores_id_tiebreak(golden):
    """rrf_fuse_topk = reference scores, re-ordered (score desc, id asc)."""
    for c in golden["rrf"]:
        if not all(isinstance(x, int) for lst in c["lists"] for x in lst):
            continue
        ids, sc = oracle.rrf_fuse_topk([np.array(l, dtype=np.int64) for l in c["lists"]],
                                       rank_const=c["rank_const"], min_score=c["min_score"])
        ref = dict(zip(c["ids"], c["scores"]))
        assert sorted(ids.tolist()) == sorted(c["ids"])
        assert [ref[i] for i in ids.tolist()] == sc.tolist()
        keys = [(-s, i) for i, s in zip(ids.tolist(), sc.tolist())]
        assert keys == sorted(keys)


def test_cosine_and_normalize(golden):
    for c in golden["cosine"]:
        assert float(oracle.cosine_similarity_pair(c["a"], c["b"])) == c["out"]
    for c in golden["normalize_l2"]:
        assert [float(x) for x in oracle.normalize_l2(c["v"])] == c["out"]


def test_union_merge_matches_reference(golden):
    for c in golden["merge_union"]:
        i = c["inputs"]
        lists = [[_Doc(d) for d in i[k]] for k in ("v", "f", "g", "s", "vi")]
        out = oracle.union_merge(lists, deduplicate=c["deduplicate"])
        assert [d.to_json() for d in out] == c["docs"]
    assert golden["merge_unknown_strategy"]["message"] == "Unknown merge strategy: rrf"
    with pytest.raises(ValueError, match="Unknown merge strategy: rrf"):
        oracle.union_merge([[]], merge_strategy="rrf")


def test_rerank_fallback_matches_reference(golden):
    for c in golden["rerank_fallback"]:
        out = oracle.rerank_fallback([_Doc(d) for d in c["docs"]])
        assert [d.to_json() for d in out] == c["out"]


def test_exact_topk_against_f64_and_pairwise():
    rng = np.random.default_rng(7)
    C = rng.standard_normal((3000, 64)).astype(np.float32)
    C[17] = 0.0                      # zero-norm row -> score 0 (search_utils.py:63-64)
    C[2000] = C[5]                   # exact duplicate -> tie broken by id
    Q = np.concatenate([rng.standard_normal((6, 64)).astype(np.float32), C[5:6] * 2.0,
                        np.zeros((1, 64), np.float32)])
    s, i = oracle.exact_topk(C, Q, 7, "cosine", block_rows=1024)
    s64 = oracle.exact_topk_scores_f64(C, Q, "cosine")
    for b in range(Q.shape[0]):
        order = np.lexsort((np.arange(C.shape[0]), -s64[b]))[:7]
        got = i[b]
        # identical ids wherever the f64 gap exceeds fp32 noise
        for j in range(7):
            if got[j] != order[j]:
                assert abs(s64[b, got[j]] - s64[b, order[j]]) < 1e-5
        assert np.allclose(s[b], s64[b, got], atol=1e-5)
        for j in range(7):
            assert abs(oracle.cosine_similarity_pair(Q[b], C[got[j]]) - s[b, j]) < 1e-5
    assert i[6, 0] == 5 and i[6, 1] == 2000          # duplicate pair, smaller id first
    assert (s[7] == 0).all() and i[7].tolist() == list(range(7))   # zero query: all 0, ids ascending
    sp, ip = oracle.exact_topk(C[:3], Q[:2], 5, "ip")
    assert (ip[:, 3:] == -1).all() and np.isinf(sp[:, 3:]).all()


def test_bf16_round():
    import torch
    x = np.random.default_rng(1).standard_normal(10000).astype(np.float32) * 100
    ref = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()
    assert (oracle.bf16_round(x) == ref).all()


def test_bm25_small_known_answer():
    # 3 docs: d0 = [0,0,1], d1 = [1], d2 = [2,2,2,2]
    tokens = np.array([0, 0, 1, 1, 2, 2, 2, 2])
    doc_ptr = np.array([0, 3, 4, 8])
    tp, pd, ptf, dl = oracle.build_postings(tokens, doc_ptr, vocab=4)
    assert tp.tolist() == [0, 1, 3, 4, 4]
    assert pd.tolist() == [0, 0, 1, 2] and ptf.tolist() == [2, 1, 1, 4] and dl.tolist() == [3, 1, 4]
    imp = oracle.bm25_impacts(tp, pd, ptf, dl)
    N, avgdl, k1, b = 3, 8 / 3, 1.2, 0.75
    def f(df, tf, d):
        idf = np.log(1 + (N - df + 0.5) / (df + 0.5))
        return idf * tf * (k1 + 1) / (tf + k1 * (1 - b + b * d / avgdl))
    want = [f(1, 2, 3), f(2, 1, 3), f(2, 1, 1), f(1, 4, 4)]
    assert np.allclose(imp, want, rtol=1e-6)
    s, i = oracle.bm25_topk(tp, pd, imp, 3, [[1, 0, 1, -1], [3], [2, 1]], k=2)
    assert i[0].tolist() == [0, 1]
    assert s[0, 0] == np.float32(imp[0] + imp[1])     # term 0 then term 1
    assert i[1].tolist() == [-1, -1]
    assert i[2].tolist() == [2, 1]


def test_merge_shard_candidates():
    s = np.array([[[0.9, 0.5, 0.1]], [[0.9, 0.6, -1.0]]], dtype=np.float32)
    i = np.array([[[4, 1, 7]], [[2, 9, -1]]], dtype=np.int64)
    ms, mi = oracle.merge_shard_candidates(s, i, 4)
    assert mi[0].tolist() == [2, 4, 9, 1]
    assert ms[0].tolist() == [np.float32(0.9), np.float32(0.9), np.float32(0.6), np.float32(0.5)]


def test_mmr_matches_reference(mmr_golden):
    """oracle.mmr_rerank == maximal_marginal_relevance (search_utils.py:1884-1922) on the reference's own outputs."""
    assert len(mmr_golden) >= 10
    for c in mmr_golden:
        order, scores = oracle.mmr_rerank(np.array(c["query"]), np.array(c["candidates"]), c["mmr_lambda"], c["min_score"])
        assert order.tolist() == c["order"]
        np.testing.assert_allclose(scores, np.array(c["scores"]), rtol=0, atol=1e-12)


def test_cosine_topk_oracle_is_pinned_to_reference_ranking(topk_golden):
    """oracle.exact_topk / exact_topk_scores_f64 against FULL RANKED LISTS produced by the reference's in-tree brute-force
    cosine ranking (Neptune branch, search_utils.py:337-366 calling calculate_cosine_similarity :56-67): every pair score,
    the ranked top-20 with duplicates (ties -> lower row first), zero rows, a zero query, near-ties. This is what moves the
    cosine top-k oracle from "parity unpinned" to "pinned"."""
    from parity import check_topk_against_full_scores
    C, cases, k = topk_golden
    assert C.shape == (2000, 64) and (oracle.bf16_round(C) == C).all()          # exact in bf16: both sides see the same numbers
    Q = np.stack([c["query"] for c in cases])
    ref_full = np.stack([c["scores"] for c in cases])
    # 1. the float64 scorer reproduces the reference function's values (same formula, float64)
    full = oracle.exact_topk_scores_f64(C, Q, "cosine")
    assert np.abs(full - ref_full).max() < 1e-14
    assert (full[:, [17, 1999]] == 0).all() and (full[3] == 0).all()            # zero rows, zero query -> 0 (:63-64)
    # 2. the fp32 top-k oracle returns the reference's ranked list: same ids position by position wherever the
    #    reference's own scores separate neighbours by more than fp32 noise, same scores to fp32 rounding
    s, i = oracle.exact_topk(C, Q, k, "cosine")
    exact = 0
    for qi, c in enumerate(cases):
        if c["min_score"] > -1.5:
            continue                                                            # the score filter is exercised below
        assert np.abs(s[qi] - c["top_scores"]).max() < 5e-7
        check_topk_against_full_scores(s[qi:qi + 1], i[qi:qi + 1], ref_full[qi:qi + 1], k, tol=5e-7)
        exact += int((i[qi] == c["top_ids"]).all())
    assert exact >= 9, exact                     # every query whose top-20 has no near-tie matches id for id
    # duplicates: identical rows score identically and the lower row comes first, in the reference list and in ours
    assert cases[0]["top_ids"][:2].tolist() == [5, 1200] and i[0, :2].tolist() == [5, 1200]
    assert cases[1]["top_ids"][:3].tolist() == [300, 1500, 1900] and i[1, :3].tolist() == [300, 1500, 1900]
    assert i[3].tolist() == list(range(k)) and cases[3]["top_ids"].tolist() == list(range(k))   # zero query: row order
    # 3. the branch's `score > min_score` filter == cutting our ranked list at the threshold
    for qi, c in enumerate(cases):
        if c["min_score"] <= -1.5:
            continue
        kept = [(int(r), float(x)) for r, x in zip(i[qi], s[qi]) if x > c["min_score"]]
        assert [r for r, _ in kept] == c["top_ids"].tolist()[:len(kept)] and len(kept) == len(c["top_ids"])
    # 4. the plain-C restatement agrees with the numpy one on the same inputs
    from oracle import c_oracle
    cs_, ci_ = c_oracle.topk(c_oracle.scores(C, Q, "cosine"), k)
    assert (ci_ == i).mean() > 0.97 and np.abs(cs_ - s).max() < 1e-6
    check_topk_against_full_scores(cs_, ci_, ref_full, k, tol=5e-7)

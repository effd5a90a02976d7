"""K5 MMR rerank vs the reference's own outputs (tests/golden/mmr_golden.json: inputs are bf16-exact) and vs the
oracle on larger random cases gathered from a corpus by id. Tolerance: 1e-3 absolute on scores (north-star
floating-point tolerance; fp32 accumulation of D <= 1024 products of unit vectors is ~1e-6), order identical
wherever neighbouring reference scores are more than the tolerance apart, ties in candidate order."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _check(got_ids, got_scores, want_order, want_scores, ids_row):
    n = len(want_order)
    got_ids, got_scores = np.asarray(got_ids), np.asarray(got_scores, dtype=np.float64)
    live = got_ids >= 0
    # survivors of min_score may differ only for scores within TOL of the cut
    assert abs(int(live.sum()) - n) <= int(np.sum(np.abs(np.asarray(want_scores) - want_scores[-1]) < TOL)) if n else live.sum() <= 1
    k = min(n, int(live.sum()))
    assert (np.diff(got_scores[:k]) <= 0).all()
    pos = {int(i): p for p, i in enumerate(ids_row)}
    want = {int(o): s for o, s in zip(want_order, want_scores)}
    for j in range(k):
        p = pos[int(got_ids[j])]
        if p in want:
            assert abs(got_scores[j] - want[p]) < TOL, (j, p, got_scores[j], want[p])
    for j in range(k):
        lo_gap = j == 0 or want_scores[j - 1] - want_scores[j] > 2 * TOL
        hi_gap = j == n - 1 or want_scores[j] - want_scores[j + 1] > 2 * TOL
        if lo_gap and hi_gap:
            assert pos[int(got_ids[j])] == int(want_order[j]), (j, got_ids[j], want_order[j])


def test_mmr_matches_reference_golden(mmr_golden):
    import torch
    from super_rag_b200 import ops
    dev = torch.device("cuda:0")
    for c in mmr_golden:
        C = np.array(c["candidates"], dtype=np.float32)
        q = np.array(c["query"], dtype=np.float32)[None]
        assert (oracle.bf16_round(C) == C).all() and (oracle.bf16_round(q) == q).all()   # inputs are bf16-exact
        m = C.shape[0]
        cb, cs = ops.ingest_rows(torch.from_numpy(C).to(dev), "cosine")
        qb, _ = ops.ingest_rows(torch.from_numpy(q).to(dev), "ip")
        ones = torch.ones(1, device=dev)                                                # the reference does not normalise q
        ids = torch.arange(m, device=dev, dtype=torch.int64)[None]
        gi, gs = ops.mmr_rerank(cb, cs, qb, ones, ids, c["mmr_lambda"], c["min_score"])
        _check(gi[0].cpu().numpy(), gs[0].cpu().numpy(), c["order"], c["scores"], np.arange(m))
        if m >= 5 and c["min_score"] <= -2.0:
            order = gi[0].cpu().tolist()
            assert order.index(1) < order.index(3)        # identical embeddings: equal scores, candidate order kept


def test_mmr_gathered_candidates_vs_oracle():
    import torch
    from super_rag_b200 import ops
    rng = np.random.default_rng(12)
    dev = torch.device("cuda:0")
    n, d = 5000, 1024
    C = rng.standard_normal((n, d)).astype(np.float32)
    C[100:140] = C[100] + 0.2 * rng.standard_normal((40, d)).astype(np.float32)          # a cluster of near-duplicates
    cb, cs = ops.ingest_rows(torch.from_numpy(C).to(dev), "cosine")
    cs[777] = float("nan")                                                                # tombstone
    C_r = oracle.bf16_round(C)
    for m, B, lam in [(200, 9, 0.5), (256, 3, 0.8), (129, 4, 0.2), (40, 5, 0.5), (17, 6, 0.5)]:
        Q = rng.standard_normal((B, d)).astype(np.float32)
        qb, qs = ops.ingest_rows(torch.from_numpy(Q).to(dev), "cosine")
        Q_r = oracle.bf16_round(Q)
        ids = np.stack([rng.choice(n, m, replace=False) for _ in range(B)]).astype(np.int64)
        ids[:, 0] = 100 + np.arange(B)                # members of the cluster
        ids[:, 1] = 120 + np.arange(B)
        ids[0, 2] = 777                               # deleted row: skipped
        ids[1, 5] = -1                                # hole in the list: skipped
        id_base = 1_000_000
        gi, gs = ops.mmr_rerank(cb, cs, qb, qs, torch.from_numpy(ids + id_base * (ids >= 0)).to(dev), lam, -2.0,
                                id_base=id_base)
        gi, gs = gi.cpu().numpy(), gs.cpu().numpy()
        for b in range(B):
            keep = np.array([i >= 0 and i != 777 for i in ids[b]])
            qn = Q_r[b].astype(np.float64) / np.linalg.norm(Q_r[b].astype(np.float64))
            order, scores = oracle.mmr_rerank(qn, C_r[ids[b][keep]], lam, -2.0)
            kept_pos = np.nonzero(keep)[0]
            assert int((gi[b] >= 0).sum()) == int(keep.sum())
            _check(gi[b], gs[b], kept_pos[order], scores, ids[b] + id_base * (ids[b] >= 0))


@pytest.mark.parametrize("d", [200, 64, 1536])
def test_mmr_tensor_core_path_ragged_dims_and_sizes(d):
    """The tcgen05 Gram kernel stages 64 dimensions per k-block: a dim that is not a multiple of 64 (zero-filled tail), one
    k-block only, more k-blocks than stages; m at both sides of the 128-row split and of the CUDA-core cut (m <= 32)."""
    import torch
    from super_rag_b200 import ops
    rng = np.random.default_rng(d)
    dev = torch.device("cuda:0")
    n = 3000
    C = rng.standard_normal((n, d)).astype(np.float32)
    C[10:30] = C[10] + 0.3 * rng.standard_normal((20, d)).astype(np.float32)
    cb, cs = ops.ingest_rows(torch.from_numpy(C).to(dev), "cosine")
    C_r = oracle.bf16_round(C)
    for m, B in [(33, 5), (128, 4), (130, 3), (255, 2), (256, 2)]:
        Q = rng.standard_normal((B, d)).astype(np.float32)
        qb, qs = ops.ingest_rows(torch.from_numpy(Q).to(dev), "cosine")
        Q_r = oracle.bf16_round(Q)
        ids = np.stack([rng.choice(n, m, replace=False) for _ in range(B)]).astype(np.int64)
        ids[:, m - 1] = 10 + np.arange(B)             # cluster members at the ends of the list
        ids[:, 0] = 20 + np.arange(B)
        gi, gs = ops.mmr_rerank(cb, cs, qb, qs, torch.from_numpy(ids).to(dev), 0.6, -2.0)
        gi, gs = gi.cpu().numpy(), gs.cpu().numpy()
        for b in range(B):
            qn = Q_r[b].astype(np.float64) / np.linalg.norm(Q_r[b].astype(np.float64))
            order, scores = oracle.mmr_rerank(qn, C_r[ids[b]], 0.6, -2.0)
            assert int((gi[b] >= 0).sum()) == m
            _check(gi[b], gs[b], order, scores, ids[b])

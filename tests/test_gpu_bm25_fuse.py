"""K2 (BM25), K3 (fusion) and K4 (shard merge) parity on the GPU through the C ABI. Bit-exact."""
import json
import os

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def zipf_corpus(n_docs, vocab, mean_len, seed, s=1.0):
    rng = np.random.default_rng(seed)
    dl = np.clip(np.round(rng.normal(mean_len, mean_len / 10, n_docs)), 1, 2 * mean_len).astype(np.int64)
    p = 1.0 / np.arange(1, vocab + 1) ** s
    p /= p.sum()
    tokens = rng.choice(vocab, size=int(dl.sum()), p=p).astype(np.int64)
    doc_ptr = np.zeros(n_docs + 1, dtype=np.int64)
    np.cumsum(dl, out=doc_ptr[1:])
    return tokens, doc_ptr


def _build(tokens, doc_ptr, vocab):
    import torch
    from super_rag_b200.bm25 import Bm25Index
    idx = Bm25Index.from_tokens(torch.from_numpy(tokens).cuda(), torch.from_numpy(doc_ptr).cuda(), vocab)
    tp, pd, ptf, dl = oracle.build_postings(tokens, doc_ptr, vocab)
    imp = oracle.bm25_impacts(tp, pd, ptf, dl)
    return idx, (tp, pd, ptf, dl, imp)


@pytest.mark.parametrize("n_docs,vocab,mean_len,k", [(50_000, 5000, 100, 100), (20_000, 300, 40, 10), (3000, 50, 20, 256)])
def test_bm25_bit_exact(n_docs, vocab, mean_len, k):
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, mean_len, seed=n_docs)
    idx, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    assert (idx.term_ptr.cpu().numpy() == tp).all()
    cd, ci = idx.canonical_postings()  # the index keeps every (term, doc-block) slice in bank order
    assert (cd.cpu().numpy() == pd).all()
    assert (ci.cpu().numpy().view(np.uint32) == imp.view(np.uint32)).all(), "impacts not bit-exact"
    rng = np.random.default_rng(4)
    queries = [rng.integers(0, vocab, size=int(rng.integers(1, 9))).tolist() for _ in range(40)]
    queries += [[0], [vocab - 1], [], [0, 1, 2, 3, 4, 5, 6, 7], [5, 5, 5, -3, vocab + 10]]
    s, i = idx.search(queries, k)
    os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, queries, k)
    assert (i.cpu().numpy() == oi).all()
    assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all()


def test_bm25_massive_ties_lowest_doc_id_wins():
    # every doc is the same single token -> identical impacts everywhere
    n_docs, vocab = 40_000, 8
    tokens = np.zeros(n_docs * 3, dtype=np.int64)
    tokens[1::3] = 1
    doc_ptr = np.arange(0, 3 * n_docs + 1, 3, dtype=np.int64)
    idx, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    s, i = idx.search([[0], [0, 1], [2]], 100)
    os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, [[0], [0, 1], [2]], 100)
    assert (i.cpu().numpy() == oi).all() and i[0].tolist() == list(range(100))
    assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all()
    assert (i[2].cpu().numpy() == -1).all()


def test_bm25_shard_uses_global_stats():
    """Doc-range shards with GLOBAL df/N/avgdl give the single-index scores bit for bit."""
    import torch
    from super_rag_b200.bm25 import Bm25Index, Bm25Stats
    from super_rag_b200 import ops
    n_docs, vocab = 30_000, 1000
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, 60, seed=77)
    full, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    stats = Bm25Stats(n_docs=n_docs, df=np.diff(tp), avgdl=float(oracle.bm25.bm25_avgdl(dl)))
    queries = [[3, 17, 256], [0, 1], [999, 500, 12, 7]]
    k = 20
    parts_s, parts_i = [], []
    half = n_docs // 2
    for lo, hi in [(0, half), (half, n_docs)]:
        t = torch.from_numpy(tokens[doc_ptr[lo]:doc_ptr[hi]]).cuda()
        dp = torch.from_numpy(doc_ptr[lo:hi + 1] - doc_ptr[lo]).cuda()
        shard = Bm25Index.from_tokens(t, dp, vocab, stats=stats, id_base=lo)
        s, i = shard.search(queries, k)
        parts_s.append(s)
        parts_i.append(i)
    ms, mi = ops.merge_candidates(torch.stack(parts_s), torch.stack(parts_i), k)
    fs, fi = full.search(queries, k)
    assert (mi == fi).all() and (ms.view(torch.int32) == fs.view(torch.int32)).all()
    os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, queries, k)
    assert (mi.cpu().numpy() == oi).all()


def _pad_lists(lists, width, dtype, fill):
    out = np.full((len(lists), width), fill, dtype=dtype)
    for r, l in enumerate(lists):
        out[r, :len(l)] = l
    return out


def test_rrf_matches_reference_golden(golden):
    """K3 on the reference's own rrf() outputs (int-id cases of tests/golden/fusion_golden.json)."""
    import torch
    from super_rag_b200 import ops
    n = 0
    for c in golden["rrf"]:
        if not c["lists"] or not all(isinstance(x, int) for l in c["lists"] for x in l):
            continue
        width = max([len(l) for l in c["lists"]] + [1])
        ids = torch.from_numpy(_pad_lists(c["lists"], width, np.int64, -1)).cuda()[:, None, :].contiguous()
        k_out = max(1, len(c["ids"]) + 3)
        oi, osc = ops.fuse(ids, None, ops.FUSE_RRF, k_out, rank_const=c["rank_const"], min_score=c["min_score"])
        oi, osc = oi.cpu().numpy()[0], osc.cpu().numpy()[0]
        m = len(c["ids"])
        ref = dict(zip(c["ids"], c["scores"]))
        assert sorted(oi[:m].tolist()) == sorted(c["ids"]) and (oi[m:] == -1).all()
        assert [ref[i] for i in oi[:m].tolist()] == osc[:m].tolist()          # float64 bit-exact
        want_i, want_s = oracle.rrf_fuse_topk([np.array(l, dtype=np.int64) for l in c["lists"]],
                                              c["rank_const"], c["min_score"], k_out)
        assert oi[:m].tolist() == want_i.tolist()
        n += 1
    assert n >= 20


def test_fuse_batched_hybrid_shape():
    """C4 shape: vector top-100 + BM25 top-100 -> top-10, 64 queries, rrf (rank_const 1 and 60) and weighted."""
    import torch
    from super_rag_b200 import ops
    rng = np.random.default_rng(8)
    B, K = 64, 100
    a = np.stack([rng.choice(400, K, replace=False) for _ in range(B)]).astype(np.int64)
    b = np.stack([rng.choice(400, K, replace=False) for _ in range(B)]).astype(np.int64)
    b[:, 90:] = -1                                                 # ragged second list
    sa = np.sort(rng.random((B, K)).astype(np.float32))[:, ::-1].copy()
    sb = np.sort(rng.random((B, K)).astype(np.float32) * 20)[:, ::-1].copy()
    ids = torch.from_numpy(np.stack([a, b])).cuda()
    sc = torch.from_numpy(np.stack([sa, sb])).cuda()
    for rc in (1, 60):
        oi, osc = ops.fuse(ids, None, ops.FUSE_RRF, 10, rank_const=rc)
        for q in range(B):
            wi, ws = oracle.rrf_fuse_topk([a[q], b[q]], rc, 0.0, 10)
            assert oi[q].cpu().numpy().tolist() == wi.tolist()
            assert osc[q].cpu().numpy().tolist() == ws.tolist()
    oi, osc = ops.fuse(ids, sc, ops.FUSE_WEIGHTED, 10, min_score=float("-inf"), weights=[0.7, 0.015])
    for q in range(B):
        wi, ws = oracle.weighted_fuse_topk([a[q], b[q]], [sa[q], sb[q]], [0.7, 0.015], 10)
        assert oi[q].cpu().numpy().tolist() == wi.tolist()
        assert osc[q].cpu().numpy().astype(np.float32).tolist() == ws.tolist()


def test_merge_candidates_matches_oracle():
    import torch
    from super_rag_b200 import ops
    rng = np.random.default_rng(9)
    G, B, KK = 8, 33, 100
    s = rng.random((G, B, KK)).astype(np.float32)
    s[:, :, 50:60] = s[:, :, 40:50]                   # ties across positions
    s[1] = s[0]                                       # ties across shards
    i = rng.permutation(G * B * KK).reshape(G, B, KK).astype(np.int64)
    i[3, :, 90:] = -1
    ms, mi = ops.merge_candidates(torch.from_numpy(s).cuda(), torch.from_numpy(i).cuda(), 100)
    ws, wi = oracle.merge_shard_candidates(s, i, 100)
    assert (mi.cpu().numpy() == wi).all() and (ms.cpu().numpy() == ws).all()


def test_bm25_many_terms_per_query():
    """Queries longer than one producer batch (32 term slots) and longer than 64: summation order stays
    ascending term id across the batches."""
    n_docs, vocab = 60_000, 400
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, 60, seed=77)
    idx, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    rng = np.random.default_rng(5)
    queries = [rng.choice(vocab, size=n, replace=False).tolist() for n in (9, 31, 32, 33, 40, 64, 65, 70, 100)]
    s, i = idx.search(queries, 50)
    os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, queries, 50)
    assert (i.cpu().numpy() == oi).all()
    assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all()


def test_bm25_two_phase_with_ties_at_the_threshold():
    """43 blocks -> sampled first phase + thresholded second phase. Documents are drawn from a handful of
    shapes, so thousands of documents tie exactly at the sample's k-th best score: ties at tau are kept and
    the globally lowest doc ids win, whichever phase their block belongs to."""
    n_docs, vocab = 700_000, 6
    rng = np.random.default_rng(9)
    shapes = np.array([[0, 1, 2], [0, 0, 1], [3, 4, 5], [0, 3, 3], [1, 1, 1], [2, 4, 4]], dtype=np.int64)
    pick = rng.integers(0, len(shapes), n_docs)
    tokens = shapes[pick].reshape(-1)
    doc_ptr = np.arange(0, 3 * n_docs + 1, 3, dtype=np.int64)
    idx, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    queries = [[0], [0, 1], [3, 4, 5], [1, 2, 4], [0, 1, 2, 3, 4, 5], [5]]
    for k in (10, 100, 256):
        s, i = idx.search(queries, k)
        os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, queries, k)
        assert (i.cpu().numpy() == oi).all(), k
        assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all(), k


def test_bm25_bank_order_is_a_slice_permutation_and_cuts_conflicts(monkeypatch):
    """The searched layout: every slice starts on a group of 8 postings, is padded to whole groups with zero-impact
    postings (trash offsets), holds the same (doc, impact) pairs as the canonical slice, and — bank-ordered — is laid
    out as oracle/bank_order.py restates it."""
    import torch
    from super_rag_b200 import ops
    from super_rag_b200.bm25 import Bm25Index
    n_docs, vocab = 70_000, 60
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, 60, seed=77)
    monkeypatch.setenv("SRAG_BM25_BANK_ORDER", "0")
    plain = Bm25Index.from_tokens(torch.from_numpy(tokens).cuda(), torch.from_numpy(doc_ptr).cuda(), vocab)
    monkeypatch.delenv("SRAG_BM25_BANK_ORDER")
    idx = Bm25Index.from_tokens(torch.from_numpy(tokens).cuda(), torch.from_numpy(doc_ptr).cuda(), vocab)
    assert not plain.bank_ordered and idx.bank_ordered
    bd = ops.bm25_block_docs()
    tp, pd, ptf, dl = oracle.build_postings(tokens, doc_ptr, vocab)
    imp = oracle.bm25_impacts(tp, pd, ptf, dl).view(np.uint32)
    d0, w0 = plain.post_doc.cpu().numpy(), plain.post_imp.cpu().numpy().view(np.uint32)
    d1, w1 = idx.post_doc.cpu().numpy(), idx.post_imp.cpu().numpy().view(np.uint32)
    sp = idx.slice_ptr.cpu().numpy().view(np.uint32).astype(np.int64) * 8
    assert (sp == plain.slice_ptr.cpu().numpy().view(np.uint32).astype(np.int64) * 8).all()
    assert len(d1) == sp[-1, -1] and len(d1) % 8 == 0
    # every slice: same pairs as the canonical postings of (term, sub-block), all inside the sub-block, padding = (-1, 0);
    # the layout itself is the one oracle/bank_order.py restates (checked on a sample of slices: it is a Python loop)
    from oracle import bank_order as bo
    t0 = t1 = 0
    exact = 0
    n_sub = sp.shape[1] - 1
    for t in range(vocab):
        docs_t, imps_t = pd[tp[t]:tp[t + 1]], imp[tp[t]:tp[t + 1]]
        for j in range(n_sub):
            lo, hi = int(sp[t, j]), int(sp[t, j + 1])
            sel = (docs_t // bd) == j
            a, wa = docs_t[sel], imps_t[sel]
            n = len(a)
            assert hi - lo == (n + 7) // 8 * 8, (t, j)
            if n == 0:
                continue
            b, wb = d1[lo:lo + n], w1[lo:lo + n]
            assert (d1[lo + n:hi] == -1).all() and (w1[lo + n:hi] == 0).all() and (d0[lo + n:hi] == -1).all()
            assert (d0[lo:lo + n] == a).all() and (w0[lo:lo + n] == wa).all()          # SRAG_BM25_BANK_ORDER=0: ascending
            o = np.argsort(b, kind="stable")
            assert (b[o] == a).all() and (wb[o] == wa).all()
            if n >= 64:
                t0 += bo.access_wavefronts(a, 0)[0]
                t1 += bo.access_wavefronts(b, 0)[0]
            if exact < 60 and (n >= 40 or (n >= 3 and exact % 3 == 0)):
                wd, ww = bo.bank_order_slice(a, wa, 0)
                assert (wd == b).all() and (ww == wb).all(), (t, j)
                exact += 1
    assert exact >= 30
    assert t1 < 0.8 * t0, (t0, t1)
    # the raw trash offsets: sub-block 0 of a block -> byte 8192, sub-block m -> m * 8192 - 4
    off = idx.post_off.cpu().numpy().view(np.uint16)
    for t in (0, vocab - 1):
        for j in range(min(n_sub, 12)):
            lo, hi = int(sp[t, j]), int(sp[t, j + 1])
            pads = off[lo:hi][w1[lo:hi] == 0]
            m = j % 8
            assert (pads == (8192 if m == 0 else m * 8192 - 4)).all()
    # same scores and ids, bit for bit, from both layouts
    rng = np.random.default_rng(5)
    queries = [rng.integers(0, vocab, size=int(rng.integers(1, 9))).tolist() for _ in range(64)]
    s0, i0 = plain.search(queries, 50)
    s1, i1 = idx.search(queries, 50)
    assert torch.equal(i0, i1) and torch.equal(s0.view(torch.int32), s1.view(torch.int32))
    # deterministic: a second build gives the same layout
    again = Bm25Index.from_tokens(torch.from_numpy(tokens).cuda(), torch.from_numpy(doc_ptr).cuda(), vocab)
    assert torch.equal(again.post_off, idx.post_off) and torch.equal(again.post_imp.view(torch.int32), idx.post_imp.view(torch.int32))


def test_bm25_sub_kernel_candidate_paths_and_fallback():
    """Phase 1 runs one warp per (query, 2048-document sub-block) and appends the documents that cross tau to the
    query's candidate region. Exercised here, bit-exact against the oracle:
      * a term the block sample never sees (tau = 0: every matched document is a candidate) whose documents are spread
        out — the append path with tau = 0;
      * the same with ~1000 documents packed into two sub-blocks: more than 64 crossings per sub-block flag the query,
        its region is cleared and it is redone by the exact per-block kernel;
      * a frequent term with graded scores (tau > 0, ordinary crossing candidates), alone and mixed with the others;
      * a document count that is not a multiple of the sub-block."""
    n_docs, vocab = 600_123, 5
    rng = np.random.default_rng(21)
    filler = 1 + (np.arange(n_docs) % 5)                     # term 3 repeated: document lengths 1..5 (+ the others)
    has1 = (np.arange(n_docs) % 7) == 0
    tf1 = np.where(has1, 1 + (np.arange(n_docs) % 3), 0)
    has0 = (np.arange(n_docs) >= 20_000) & (np.arange(n_docs) < 21_000)
    rare = np.zeros(n_docs, dtype=bool)
    rare[rng.choice(n_docs, 50, replace=False)] = True
    rare[:16384] = False                                     # keep it out of sampled block 0
    counts = np.stack([has0.astype(np.int64), tf1.astype(np.int64), rare.astype(np.int64), filler.astype(np.int64)], 1)
    dl = counts.sum(1)
    doc_ptr = np.zeros(n_docs + 1, dtype=np.int64)
    np.cumsum(dl, out=doc_ptr[1:])
    tokens = np.repeat(np.tile(np.arange(4, dtype=np.int64), n_docs), counts.reshape(-1))
    idx, (tp, pd, ptf, dl_o, imp) = _build(tokens, doc_ptr, vocab)
    queries = [[0], [2], [1], [1, 2], [0, 1], [0, 1, 2], [0, 2, 3], [4]]
    for k in (10, 100):
        s, i = idx.search(queries, k)
        os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, queries, k)
        assert (i.cpu().numpy() == oi).all(), k
        assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all(), k


def test_bm25_native_build_matches_oracle_postings():
    """srag_bm25_build_plan / _fill (per-document sort + run-length encode, counting, scans, scatter, per-slice sort)
    against oracle.build_postings: term_ptr, postings (documents ascending inside a term), tf, dl and the sub-block
    offsets, all exact. Covers empty documents, a single-token document repeated 5000 times, a 20 000-token document (long-document pass), slices longer than 32
    postings (bitonic path) and short ones (ranking path), and the two error flags."""
    import torch
    from super_rag_b200 import ops
    from super_rag_b200._lib import SragError
    n_docs, vocab = 9000, 700
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, 60, seed=3)
    # splice in: an empty document, a very long repetitive one, a long varied one
    dl = np.diff(doc_ptr)
    docs = [tokens[doc_ptr[d]:doc_ptr[d + 1]] for d in range(n_docs)]
    docs[17] = docs[17][:0]
    docs[4000] = np.full(5000, vocab - 1, dtype=np.int64)
    docs[8999] = np.arange(8192, dtype=np.int64) % vocab
    docs[123] = (np.arange(20_000, dtype=np.int64) * 7) % vocab          # longer than the main pass takes: long-document pass
    tokens = np.concatenate(docs)
    doc_ptr = np.zeros(n_docs + 1, dtype=np.int64)
    np.cumsum([len(x) for x in docs], out=doc_ptr[1:])
    tp, pd, ptf, dl = oracle.build_postings(tokens, doc_ptr, vocab)
    t = torch.from_numpy(tokens).cuda()
    dp = torch.from_numpy(doc_ptr).cuda()
    term_ptr, post_doc, post_tf, dl_g, df, block_ptr = ops.bm25_build_from_tokens(t, dp, vocab)
    assert (term_ptr.cpu().numpy() == tp).all()
    assert (post_doc.cpu().numpy() == pd).all()
    assert (post_tf.cpu().numpy() == ptf).all()
    assert (dl_g.cpu().numpy() == dl).all()
    assert (df.cpu().numpy() == np.diff(tp)).all()
    assert torch.equal(block_ptr, ops.bm25_block_index(term_ptr, post_doc, n_docs))
    assert (t.cpu().numpy() == tokens).all(), "the caller's tokens must not be consumed"
    # deterministic
    again = ops.bm25_build_from_tokens(t, dp, vocab)
    assert torch.equal(again[1], post_doc) and torch.equal(again[2], post_tf)
    # errors are loud
    bad = t.clone()
    bad[5] = vocab
    with pytest.raises(SragError):
        ops.bm25_build_from_tokens(bad, dp, vocab)
    long_ptr = torch.tensor([0, 57345], dtype=torch.int64, device="cuda")
    with pytest.raises(SragError):
        ops.bm25_build_from_tokens(torch.zeros(57345, dtype=torch.int64, device="cuda"), long_ptr, vocab)
    # degenerate: no documents
    e = ops.bm25_build_from_tokens(torch.zeros(0, dtype=torch.int64, device="cuda"),
                                   torch.zeros(1, dtype=torch.int64, device="cuda"), 5)
    assert e[0].tolist() == [0] * 6 and e[1].numel() == 0


def test_bm25_sub_kernel_queries_longer_than_32_terms():
    """Two-phase path (35 blocks) with queries of 33 .. 90 terms: the warp-granular kernel walks the term slots in
    groups of 32; summation order must stay ascending term id across the groups."""
    n_docs, vocab = 560_000, 200
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, 12, seed=5)
    idx, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    rng = np.random.default_rng(6)
    queries = [rng.choice(vocab, size=n, replace=False).tolist() for n in (33, 40, 64, 65, 90, 8)]
    s, i = idx.search(queries, 50)
    os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, queries, 50)
    assert (i.cpu().numpy() == oi).all()
    assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all()


@pytest.mark.parametrize("filtered", [False, True])
def test_bm25_two_phase_single_query_odd_k(filtered):
    """B = 1 with k in {1, 5} on an index of >= 32 blocks (the default fulltext_search node shape: one query, top_k = 5).
    B * k is odd, so the sample's (ids, scores) region of the workspace must keep its 8-byte id stores aligned for any
    B * k (round-1 layout put the ids behind B * k floats: a misaligned-address fault on exactly this shape)."""
    import torch
    n_docs, vocab = 560_000, 200
    tokens, doc_ptr = zipf_corpus(n_docs, vocab, 12, seed=5)
    idx, (tp, pd, ptf, dl, imp) = _build(tokens, doc_ptr, vocab)
    kw = {}
    valid = None
    if filtered:
        tags = (1 << (np.arange(n_docs) % 3)).astype(np.int32)
        kw = {"doc_tags": torch.from_numpy(tags).cuda(), "q_tag_mask": torch.tensor([0b101], dtype=torch.int32).cuda()}
        valid = (tags & 0b101) != 0
    for k in (1, 5, 7):
        for q in ([3, 17, 150], [0], [199, 42]):
            s, i = idx.search([q], k, **kw)
            torch.cuda.synchronize()
            os_, oi = oracle.bm25_topk(tp, pd, imp, n_docs, [q], k, valid=valid)
            assert (i.cpu().numpy() == oi).all(), (k, q)
            assert (s.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all(), (k, q)


def test_fulltext_collection_whose_texts_tokenise_to_nothing():
    """avgdl = 0 and nnz = 0: the index builds and every search returns no hits (srag_bm25_impacts used to reject it)."""
    from super_rag_b200.fulltext import FulltextCollection, hash_tokenizer
    col = FulltextCollection("empty-texts", hash_tokenizer(1024), 1024, device="cuda")
    col.add(["", "   ", "!!!"])
    assert col.search("anything at all", 5) == []
    col.add(["now a real word"])
    assert [d.text for d in col.search("word", 5)] == ["now a real word"]


def test_exchange_keys_pack_merge_and_fused_emission():
    """The 8-byte cross-shard exchange format: srag_pack_keys == the torch statement (sharded.pack_keys_torch);
    srag_merge_keys == srag_merge_candidates == the oracle on ties across positions and shards, ragged lists, -0.0 and
    negative scores; srag_topk_search_keys emits exactly pack(scores, ids) of the ordinary search."""
    import torch
    from super_rag_b200 import ops
    from super_rag_b200.sharded import pack_keys_torch, unpack_keys_torch
    rng = np.random.default_rng(19)
    G, B, KK = 8, 37, 100
    s = rng.standard_normal((G, B, KK)).astype(np.float32)
    s[:, :, 50:60] = s[:, :, 40:50]
    s[1] = s[0]
    s[2, :, :5] = -0.0
    s[3, :, :5] = 0.0
    i = rng.permutation(G * B * KK).reshape(G, B, KK).astype(np.int64) + 4_000_000_000 - G * B * KK - 8
    i[3, :, 90:] = -1
    st, it = torch.from_numpy(s).cuda(), torch.from_numpy(i).cuda()
    keys = ops.pack_keys(st, it)
    assert torch.equal(keys.cpu(), pack_keys_torch(torch.from_numpy(s), torch.from_numpy(i)))
    us, ui = unpack_keys_torch(keys)
    m = it >= 0
    assert torch.equal(ui[m], it[m]) and torch.equal(us[m], st[m]) and (ui[~m] == -1).all()       # -0.0 comes back as +0.0
    for k in (1, 10, 100):
        ms, mi = ops.merge_keys(keys, k)
        cs, ci = ops.merge_candidates(st, it, k)
        ws, wi = oracle.merge_shard_candidates(s, i, k)
        assert (mi.cpu().numpy() == wi).all() and (ci.cpu().numpy() == wi).all(), k
        assert (ms.cpu().numpy() == ws).all() and (cs.cpu().numpy() == ws).all(), k
    # fused emission from the finalize kernel
    n, d, b, k = 30_000, 256, 200, 10
    C = torch.randn(n, d, device="cuda")
    Q = torch.randn(b, d, device="cuda")
    C[20_000] = C[5]
    Q[0] = C[5]
    cb, cs_ = ops.ingest_rows(C, "cosine")
    qb, qs = ops.ingest_rows(Q, "cosine")
    base = 1_000_000_007
    s0, i0 = ops.topk_search(cb, cs_, qb, qs, k, id_base=base)
    kbuf = torch.empty(b, k, dtype=torch.int64, device="cuda")
    s1, i1 = ops.topk_search(cb, cs_, qb, qs, k, id_base=base, out_keys=kbuf)
    assert torch.equal(i0, i1) and torch.equal(s0, s1) and torch.equal(kbuf, ops.pack_keys(s0, i0))
    kbuf2 = torch.zeros_like(kbuf)
    assert ops.topk_search(cb, cs_, qb, qs, k, id_base=base, out_keys=kbuf2, keys_only=True) == (None, None)
    assert torch.equal(kbuf2, kbuf)
    with pytest.raises(Exception):                          # global rows that do not fit 32 bits are refused loudly
        ops.topk_search(cb, cs_, qb, qs, k, id_base=1 << 32, out_keys=kbuf)


def test_fuse_at_the_abi_limit_with_duplicates():
    """K3 is a segmented sort: 8 lists x 256 entries = the ABI's 2048-entry limit, ids drawn from a small pool so that
    most chunks occur in several lists and some twice inside one list (the reference adds those again)."""
    import torch
    from super_rag_b200 import ops
    rng = np.random.default_rng(23)
    L, B, K = 8, 5, 256
    ids = rng.integers(0, 700, (L, B, K)).astype(np.int64)
    ids[:, :, 250:] = -1
    ids[1, 2, :] = -1                                              # an empty list
    sc = rng.random((L, B, K)).astype(np.float32)
    it = torch.from_numpy(ids).cuda()
    for rc in (1, 60):
        oi, osc = ops.fuse(it, None, ops.FUSE_RRF, 300, rank_const=rc, min_score=0.02 if rc == 1 else 0.0)
        for q in range(B):
            wi, ws = oracle.rrf_fuse_topk([ids[l, q] for l in range(L)], rc, 0.02 if rc == 1 else 0.0, 300)
            got_i, got_s = oi[q].cpu().numpy(), osc[q].cpu().numpy()
            assert got_i[:len(wi)].tolist() == wi.tolist() and (got_i[len(wi):] == -1).all()
            assert got_s[:len(ws)].tolist() == ws.tolist()
    w = [0.5, 0.25, 1.0, 0.125, 2.0, 0.75, 0.3, 0.9]
    oi, osc = ops.fuse(it, torch.from_numpy(sc).cuda(), ops.FUSE_WEIGHTED, 50, min_score=float("-inf"), weights=w)
    for q in range(B):
        wi, ws = oracle.weighted_fuse_topk([ids[l, q] for l in range(L)], [sc[l, q] for l in range(L)], w, 50)
        assert oi[q].cpu().numpy()[:len(wi)].tolist() == wi.tolist()
        assert osc[q].cpu().numpy()[:len(ws)].astype(np.float32).tolist() == ws.tolist()

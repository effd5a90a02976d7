#!/usr/bin/env python
"""C3 / C4 measurements (BASELINE.json configs[2], configs[3]) on one B200. Not the bench contract
(bench.py is); prints one JSON line per configuration.

  C3  BM25 over a synthetic Zipfian inverted index: N docs, V = 50k vocab, dl ~ N(400, 40), top-100
  C4  hybrid: vector top-100 (K1) + BM25 top-100 (K2) -> RRF (K3) -> top-10 over the same N ids

    python tools/bench_hybrid.py --docs 10000000 [--skip-c4] [--check 4]

The index is generated on the device in document chunks (seeded, two passes: count, then scatter into
the term-major CSR), so no 32 GB token array ever exists.

This module never imports `oracle/`: the full-size oracle checks and the CPU baselines are HOOKS that bench.py (the
one place allowed to execute the oracle next to tests/) passes to run(): for the check queries the build keeps the
(doc, tf) postings of their terms and the document lengths on the host, the hook recomputes impacts and scores with
oracle/bm25.py and compares ids and score bits. Stand-alone (`python tools/bench_hybrid.py`) the lines carry
"oracle_check": null.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402
from super_rag_b200.bm25 import Bm25Index, Bm25Stats  # noqa: E402

VOCAB = 50_000
CHUNK_DOCS = 100_000


def zipf_cdf(vocab, dev):
    p = 1.0 / torch.arange(1, vocab + 1, dtype=torch.float64, device=dev)
    return torch.cumsum(p / p.sum(), 0).to(torch.float32)


def gen_chunk(c, n_docs, cdf, dev):
    """(term int64, doc_local int64, tf int32) sorted by (term, doc) + dl int32 for documents of chunk c."""
    lo, hi = c * CHUNK_DOCS, min(n_docs, (c + 1) * CHUNK_DOCS)
    g = torch.Generator(device=dev)
    g.manual_seed(3_000_017 + c)
    m = hi - lo
    dl = torch.clamp(torch.round(400 + 40 * torch.randn(m, generator=g, device=dev)), 50, 800).to(torch.int64)
    total = int(dl.sum().item())
    u = torch.rand(total, generator=g, device=dev)
    term = torch.searchsorted(cdf, u).clamp_(max=VOCAB - 1)
    doc = torch.repeat_interleave(torch.arange(m, device=dev, dtype=torch.int64), dl)
    key, _ = torch.sort(term * m + doc)
    uniq, tf = torch.unique_consecutive(key, return_counts=True)
    t = torch.div(uniq, m, rounding_mode="floor")
    return t, uniq - t * m + lo, tf.to(torch.int32), dl.to(torch.int32)


def build_index(n_docs, dev, keep_terms=None, kept=None):
    """`keep_terms` (sorted unique term ids): their canonical (ascending document) postings (doc int32, tf int32) and the
    document lengths are copied to the host into `kept` = {"terms": {t: (doc, tf)}, "dl": int32 [n_docs]} before the
    device copy of tf is dropped — the inputs of the full-size oracle check."""
    cdf = zipf_cdf(VOCAB, dev)
    n_chunks = (n_docs + CHUNK_DOCS - 1) // CHUNK_DOCS
    counts = torch.zeros(n_chunks, VOCAB, dtype=torch.int64, device=dev)
    dl_all = torch.empty(n_docs, dtype=torch.int32, device=dev)
    for c in range(n_chunks):
        t, d, tf, dl = gen_chunk(c, n_docs, cdf, dev)
        counts[c] = torch.bincount(t, minlength=VOCAB)
        dl_all[c * CHUNK_DOCS:c * CHUNK_DOCS + dl.numel()] = dl
    df = counts.sum(0)
    term_ptr = torch.zeros(VOCAB + 1, dtype=torch.int64, device=dev)
    term_ptr[1:] = torch.cumsum(df, 0)
    nnz = int(term_ptr[-1].item())
    chunk_base = term_ptr[:-1][None, :] + torch.cumsum(counts, 0) - counts      # [chunk, term] first slot
    post_doc = torch.zeros(nnz + 8, dtype=torch.int32, device=dev)[:nnz]
    post_tf = torch.empty(nnz, dtype=torch.int32, device=dev)
    for c in range(n_chunks):
        t, d, tf, _ = gen_chunk(c, n_docs, cdf, dev)
        seg_start = torch.cumsum(counts[c], 0) - counts[c]                     # start of term t inside the chunk
        pos = chunk_base[c][t] + (torch.arange(t.numel(), device=dev) - seg_start[t])
        post_doc[pos] = d.to(torch.int32)
        post_tf[pos] = tf
    del counts, chunk_base
    total = float(dl_all.to(torch.int64).sum().item())
    stats = Bm25Stats(n_docs=n_docs, df=df.cpu().numpy().astype(np.int64), avgdl=float(np.float32(total / n_docs)))
    if keep_terms is not None and kept is not None:
        kept["dl"] = dl_all.cpu().numpy()
        kept["terms"] = {}
        for t in keep_terms:
            a, b = int(term_ptr[t].item()), int(term_ptr[t + 1].item())
            kept["terms"][int(t)] = (post_doc[a:b].cpu().numpy(), post_tf[a:b].cpu().numpy())
    idx = Bm25Index.from_postings(term_ptr, post_doc, post_tf, dl_all, stats)
    del post_tf
    torch.cuda.empty_cache()
    return idx


def make_queries(n_q, terms_per_q, dev, lo_rank=101, seed=4):
    """Distinct terms per query, Zipf(s=1) over ranks lo_rank..V (rank r == term id r-1)."""
    rng = np.random.default_rng(seed)
    r = np.arange(lo_rank, VOCAB + 1)
    p = 1.0 / r
    p /= p.sum()
    out = np.full((n_q, terms_per_q), -1, dtype=np.int32)
    for q in range(n_q):
        out[q] = np.sort(rng.choice(r, size=terms_per_q, replace=False, p=p) - 1)
    return torch.from_numpy(out).to(dev)


def torch_restatement(idx, terms, k):
    """Independent torch scorer over the CANONICAL postings (term_ptr indexes those; the searched arrays are padded and
    bank-ordered): index_add in ascending term order == the kernel's summation order."""
    if getattr(idx, "_canon_cache", None) is None:
        idx._canon_cache = idx.canonical_postings()
    post_doc, post_imp = idx._canon_cache
    acc = torch.zeros(idx.n_docs, dtype=torch.float32, device=idx.device)
    hit = torch.zeros(idx.n_docs, dtype=torch.bool, device=idx.device)
    for t in sorted(set(int(x) for x in terms if x >= 0)):
        a, b = int(idx.term_ptr[t].item()), int(idx.term_ptr[t + 1].item())
        d = post_doc[a:b].to(torch.int64)
        acc[d] = acc[d] + post_imp[a:b]              # unique docs inside a term: one add per doc, fixed order
        hit[d] = True
    key = torch.where(hit, acc, torch.full_like(acc, -1.0))
    s, i = torch.sort(key, descending=True, stable=True)   # stable: equal scores keep ascending doc id
    return s[:k], i[:k]


def timed(fn, warmup, steps):
    # warm up for at least `warmup` calls AND 0.3 s: after seconds of host-only work (index build bookkeeping, the oracle
    # hooks) the GPU sits in a low power state and the first calls run at idle clocks (one bench run showed C4 at 37.5 ms
    # against 27.8 ms measured right after other GPU work)
    t0 = time.perf_counter()
    n = 0
    while n < warmup or time.perf_counter() - t0 < 0.3:
        fn()
        n += 1
        if n >= warmup:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=1024)
    ap.add_argument("--terms", type=int, default=8)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--check", type=int, default=16)
    ap.add_argument("--skip-c4", action="store_true")
    ap.add_argument("--skip-stopwords", action="store_true")
    ap.add_argument("--dim", type=int, default=1024)
    args = ap.parse_args()
    for line in run(args):
        print(json.dumps(line), flush=True)


class Hooks:
    """What bench.py plugs in (it may execute oracle/; this module may not). Every hook gets the context dict and
    returns a JSON-able dict (or None)."""
    c3_check = None      # full-size oracle check of the C3 results
    c3_cpu = None        # CPU baseline for C3
    c4_check = None
    c4_cpu = None


def run(args, dev=None, hooks=None):
    """Generator of the C3 line, the stop-word C3 line and (unless args.skip_c4) the C4 line; bench.py attaches them to
    its own line."""
    dev = dev or torch.device("cuda:0")
    hooks = hooks or Hooks()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    hbm_src = "measured" if "hbm_gbs" in peaks else "fallback"

    qt = make_queries(args.queries, args.terms, dev)
    n_check = max(0, min(int(getattr(args, "check", 0) or 0), args.queries))
    # the stop-word variant (SURVEY §8d "a second run includes them"): ranks 1..V, df up to ~N per term
    qt_stop = make_queries(args.queries, args.terms, dev, lo_rank=1, seed=5)
    n_check_stop = min(4, n_check)
    keep = sorted(set(int(t) for t in qt[:n_check].reshape(-1).tolist() if t >= 0) |
                  set(int(t) for t in qt_stop[:n_check_stop].reshape(-1).tolist() if t >= 0))
    kept = {}
    t0 = time.time()
    idx = build_index(args.docs, dev, keep_terms=keep if n_check else None, kept=kept)
    torch.cuda.synchronize()
    build_s = time.time() - t0
    post_bytes = idx.query_posting_bytes(qt)
    ms, (bs, bi) = timed(lambda: idx.search_terms(qt, 100), args.warmup, args.steps)
    sorted_ok = bool((bs[:, :-1] >= bs[:, 1:]).all().item())
    # e2e through the public API with HOST buffers: pinned int32 term ids -> H2D -> K2 -> D2H of (scores, ids)
    qt_host = qt.cpu().pin_memory()
    out_s = torch.empty(args.queries, 100, dtype=torch.float32).pin_memory()
    out_i = torch.empty(args.queries, 100, dtype=torch.int64).pin_memory()

    def e2e_c3():
        s_, i_ = idx.search_terms(qt_host, 100)
        out_s.copy_(s_, non_blocking=True)
        out_i.copy_(i_, non_blocking=True)
        torch.cuda.synchronize()

    for _ in range(2):
        e2e_c3()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_c3()
    e2e_ms = (time.perf_counter() - t0) / args.steps * 1e3
    ctx = {"idx": idx, "qt": qt, "qt_stop": qt_stop, "n_check": n_check, "n_check_stop": n_check_stop, "kept": kept,
           "scores": bs, "ids": bi, "k": 100, "n_docs": args.docs, "args": args}
    line = {"config": "C3", "workload": f"BM25 {args.docs} docs, V={VOCAB}, nnz={idx.nnz}, {args.queries} queries x "
            f"{args.terms} terms (Zipf ranks 101..V), top-100", "queries_per_s": args.queries / (ms * 1e-3),
            "ms_per_batch": ms, "posting_bytes_per_batch": post_bytes,
            "roofline": {"bound": "hbm", "achieved": post_bytes / (ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": post_bytes / (ms * 1e-3) / 1e9 / hbm, "peak_source": hbm_src,
                         "algorithmic_bytes": "8 B x sum of df over the query terms (SURVEY 8d)",
                         "index_bytes_per_posting": idx.bytes_per_posting() if hasattr(idx, "bytes_per_posting") else 8,
                         "traffic": bm25_traffic(args.docs, args.queries), "traffic_source": "ncu capture under profiles/ (see profiles/bm25_traffic.json), not this run"},
            "e2e": {"value": args.queries / (e2e_ms * 1e-3), "unit": "queries/s", "ms_per_batch": e2e_ms,
                    "h2d_bytes_per_step": int(qt_host.numel() * 4), "d2h_bytes_per_step": args.queries * 100 * 12},
            "index_bytes": idx.nnz_stored * idx.bytes_per_posting(), "index_padding_postings": idx.nnz_stored - idx.nnz, "index_build_s": build_s, "scores_sorted": sorted_ok,
            "oracle_check": hooks.c3_check(ctx) if hooks.c3_check and n_check else None,
            "cpu_baseline": hooks.c3_cpu(ctx) if hooks.c3_cpu and n_check else None}
    yield line

    if not getattr(args, "skip_stopwords", False):
        pb = idx.query_posting_bytes(qt_stop)
        ms_s, (ss, si) = timed(lambda: idx.search_terms(qt_stop, 100), 1, 2)
        ctx_s = dict(ctx, qt=qt_stop, n_check=n_check_stop, scores=ss, ids=si)
        yield {"config": "C3-stopwords", "workload": f"as C3 but query terms drawn over ranks 1..V (df up to ~N per term: "
               f"the two-pass threshold and the crossing-list overflow path at full size)", "queries_per_s": args.queries / (ms_s * 1e-3),
               "ms_per_batch": ms_s, "posting_bytes_per_batch": pb,
               "roofline": {"bound": "hbm", "achieved": pb / (ms_s * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                            "frac": pb / (ms_s * 1e-3) / 1e9 / hbm},
               "oracle_check": hooks.c3_check(ctx_s) if hooks.c3_check and n_check_stop else None}
    if args.skip_c4:
        return

    g = torch.Generator(device=dev)
    g.manual_seed(1)
    corpus = torch.empty(args.docs, args.dim, dtype=torch.bfloat16, device=dev)
    for r0 in range(0, args.docs, 1 << 20):
        r1 = min(args.docs, r0 + (1 << 20))
        corpus[r0:r1] = torch.randn(r1 - r0, args.dim, generator=g, device=dev).to(torch.bfloat16)
    _, cs = ops.ingest_rows(corpus, "cosine")
    q32 = torch.randn(args.queries, args.dim, generator=g, device=dev)
    h = args.queries // 2
    src = torch.randint(0, args.docs, (h,), generator=g, device=dev)
    q32[:h] = corpus[src].float() + 0.1 * q32[:h]
    ws = torch.empty(ops.topk_workspace_bytes(args.docs, args.queries, 100, args.dim), dtype=torch.uint8, device=dev)

    def hybrid():
        qb, qs = ops.ingest_rows(q32, "cosine")
        vs, vi = ops.topk_search(corpus, cs, qb, qs, 100, workspace=ws)
        s2, i2 = idx.search_terms(qt, 100)
        return ops.fuse(torch.stack([vi, i2]), None, ops.FUSE_RRF, 10, rank_const=60), (vs, vi)

    ms4, ((fi, fs), (vs, vi)) = timed(hybrid, args.warmup, args.steps)
    # the same loop with events between the stages (no syncs): where the time goes when the kernels run back to back
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    for it in range(args.steps):
        evs[it][0].record()
        qb_, qs_ = ops.ingest_rows(q32, "cosine")
        vs_, vi_ = ops.topk_search(corpus, cs, qb_, qs_, 100, workspace=ws)
        evs[it][1].record()
        s2_, i2_ = idx.search_terms(qt, 100)
        evs[it][2].record()
        ops.fuse(torch.stack([vi_, i2_]), None, ops.FUSE_RRF, 10, rank_const=60)
        evs[it][3].record()
    torch.cuda.synchronize()
    in_loop = [sum(evs[it][j].elapsed_time(evs[it][j + 1]) for it in range(args.steps)) / args.steps for j in range(3)]
    ms_v, _ = timed(lambda: ops.topk_search(corpus, cs, *ops.ingest_rows(q32, "cosine"), 100, workspace=ws), 1, args.steps)
    ms_f, _ = timed(lambda: ops.fuse(torch.stack([vi, bi]), None, ops.FUSE_RRF, 10, rank_const=60), 1, args.steps)
    planted_ok = bool((vi[:h, 0] == src).all().item())
    # e2e through the public API (sharded.hybrid_search on one rank) with HOST buffers
    from super_rag_b200.sharded import ShardedBm25Index, ShardedVectorIndex, hybrid_search
    vec_index = ShardedVectorIndex(corpus, cs, id_base=0)
    bm_index = ShardedBm25Index(idx)
    q_host = q32.cpu().pin_memory()
    o_i = torch.empty(args.queries, 10, dtype=torch.int64).pin_memory()
    o_s = torch.empty(args.queries, 10, dtype=torch.float64).pin_memory()

    def e2e_c4():
        i_, s_ = hybrid_search(vec_index, bm_index, q_host, qt_host, 100, 10, rank_const=60)
        o_i.copy_(i_, non_blocking=True)
        o_s.copy_(s_, non_blocking=True)
        torch.cuda.synchronize()

    for _ in range(2):
        e2e_c4()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_c4()
    e2e4_ms = (time.perf_counter() - t0) / args.steps * 1e3
    same_as_e2e = bool(torch.equal(o_i.to(dev), fi))
    ctx4 = dict(ctx, corpus=corpus, row_scale=cs, q32=q32, vec_scores=vs, vec_ids=vi, fused_ids=fi, fused_scores=fs,
                rank_const=60, k_out=10)
    yield ({"config": "C4", "workload": f"hybrid: vector top-100 ({args.docs}x{args.dim} bf16) + BM25 top-100 -> RRF(60) -> top-10, "
            f"{args.queries} queries", "queries_per_s": args.queries / (ms4 * 1e-3), "ms_per_batch": ms4,
            "ms_vector_top100": ms_v, "ms_bm25_top100": ms, "ms_fuse": ms_f,
            "ms_in_pipeline": {"vector_top100": in_loop[0], "bm25_top100": in_loop[1], "fuse": in_loop[2]},
            "vector_tflops": 2.0 * args.queries * args.docs * args.dim / (ms_v * 1e-3) / 1e12,
            "e2e": {"value": args.queries / (e2e4_ms * 1e-3), "unit": "queries/s", "ms_per_batch": e2e4_ms,
                    "h2d_bytes_per_step": int(q_host.numel() * 4 + qt_host.numel() * 4),
                    "d2h_bytes_per_step": args.queries * 10 * 16, "same_ids_as_device_run": same_as_e2e},
            "planted_top1_ok": planted_ok,
            "oracle_check": hooks.c4_check(ctx4) if hooks.c4_check and n_check else None,
            "cpu_baseline": hooks.c4_cpu(ctx4) if hooks.c4_cpu and n_check else None})


def bm25_traffic(n_docs, n_queries):
    """dram bytes per main-pass launch from the committed ncu capture of this configuration (None if there is none)."""
    try:
        with open(os.path.join(os.path.dirname(__file__), "..", "profiles", "bm25_traffic.json")) as f:
            return json.load(f).get(f"dram_bytes_per_launch_docs{n_docs}_q{n_queries}")
    except Exception:
        return None


if __name__ == "__main__":
    main()

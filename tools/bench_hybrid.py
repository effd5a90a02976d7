#!/usr/bin/env python
"""C3 / C4 measurements (BASELINE.json configs[2], configs[3]) on one B200. Not the bench contract
(bench.py is); prints one JSON line per configuration.

  C3  BM25 over a synthetic Zipfian inverted index: N docs, V = 50k vocab, dl ~ N(400, 40), top-100
  C4  hybrid: vector top-100 (K1) + BM25 top-100 (K2) -> RRF (K3) -> top-10 over the same N ids

    python tools/bench_hybrid.py --docs 10000000 [--skip-c4] [--check 4]

The index is generated on the device in document chunks (seeded, two passes: count, then scatter into
the term-major CSR), so no 32 GB token array ever exists. --check Q re-scores Q queries with an
independent torch restatement (index_add in ascending term order == the kernel's summation order) and
compares ids and score bits.
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


def build_index(n_docs, dev):
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
    post_doc = torch.zeros(nnz + 4, dtype=torch.int32, device=dev)[:nnz]
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
    idx = Bm25Index.from_postings(term_ptr, post_doc, post_tf, dl_all, stats)
    del post_tf
    torch.cuda.empty_cache()
    return idx


def make_queries(n_q, terms_per_q, dev, lo_rank=101):
    """Distinct terms per query, Zipf(s=1) over ranks lo_rank..V (rank r == term id r-1)."""
    rng = np.random.default_rng(4)
    r = np.arange(lo_rank, VOCAB + 1)
    p = 1.0 / r
    p /= p.sum()
    out = np.full((n_q, terms_per_q), -1, dtype=np.int32)
    for q in range(n_q):
        out[q] = np.sort(rng.choice(r, size=terms_per_q, replace=False, p=p) - 1)
    return torch.from_numpy(out).to(dev)


def torch_restatement(idx, terms, k):
    acc = torch.zeros(idx.n_docs, dtype=torch.float32, device=idx.device)
    hit = torch.zeros(idx.n_docs, dtype=torch.bool, device=idx.device)
    for t in sorted(set(int(x) for x in terms if x >= 0)):
        a, b = int(idx.term_ptr[t].item()), int(idx.term_ptr[t + 1].item())
        d = idx.post_doc[a:b].to(torch.int64)
        acc[d] = acc[d] + idx.post_imp[a:b]          # unique docs inside a term: one add per doc, fixed order
        hit[d] = True
    key = torch.where(hit, acc, torch.full_like(acc, -1.0))
    s, i = torch.sort(key, descending=True, stable=True)   # stable: equal scores keep ascending doc id
    return s[:k], i[:k]


def timed(fn, warmup, steps):
    for _ in range(warmup):
        fn()
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
    ap.add_argument("--check", type=int, default=4)
    ap.add_argument("--skip-c4", action="store_true")
    ap.add_argument("--dim", type=int, default=1024)
    args = ap.parse_args()
    for line in run(args):
        print(json.dumps(line), flush=True)


def run(args, dev=None):
    """Generator of the C3 line and (unless args.skip_c4) the C4 line; bench.py attaches them to its own line."""
    dev = dev or torch.device("cuda:0")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)

    t0 = time.time()
    idx = build_index(args.docs, dev)
    torch.cuda.synchronize()
    build_s = time.time() - t0
    qt = make_queries(args.queries, args.terms, dev)
    post_bytes = idx.query_posting_bytes(qt)
    ms, (bs, bi) = timed(lambda: idx.search_terms(qt, 100), args.warmup, args.steps)
    ok = None
    if args.check:
        ok = True
        for q in range(args.check):
            ws, wi = torch_restatement(idx, qt[q].tolist(), 100)
            n = int((wi >= 0).sum().item())
            ok &= bool(torch.equal(bi[q, :n], wi[:n])) and bool(torch.equal(bs[q, :n].view(torch.int32), ws[:n].view(torch.int32)))
    sorted_ok = bool((bs[:, :-1] >= bs[:, 1:]).all().item())
    yield ({"config": "C3", "workload": f"BM25 {args.docs} docs, V={VOCAB}, nnz={idx.nnz}, {args.queries} queries x "
                      f"{args.terms} terms (Zipf ranks 101..V), top-100", "queries_per_s": args.queries / (ms * 1e-3),
                      "ms_per_batch": ms, "posting_bytes_per_batch": post_bytes,
                      "roofline": {"bound": "hbm", "achieved": post_bytes / (ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                                   "frac": post_bytes / (ms * 1e-3) / 1e9 / hbm},
                      "index_bytes": idx.nnz * 8, "index_build_s": build_s, "restatement_check_ok": ok,
                      "scores_sorted": sorted_ok})
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
        return ops.fuse(torch.stack([vi, i2]), None, ops.FUSE_RRF, 10, rank_const=60), vi

    ms4, ((fi, fs), vi) = timed(hybrid, args.warmup, args.steps)
    ms_v, _ = timed(lambda: ops.topk_search(corpus, cs, *ops.ingest_rows(q32, "cosine"), 100, workspace=ws), 1, args.steps)
    ms_f, _ = timed(lambda: ops.fuse(torch.stack([vi, bi]), None, ops.FUSE_RRF, 10, rank_const=60), 1, args.steps)
    planted_ok = bool((vi[:h, 0] == src).all().item())
    yield ({"config": "C4", "workload": f"hybrid: vector top-100 ({args.docs}x{args.dim} bf16) + BM25 top-100 -> RRF(60) -> top-10, "
                      f"{args.queries} queries", "queries_per_s": args.queries / (ms4 * 1e-3), "ms_per_batch": ms4,
                      "ms_vector_top100": ms_v, "ms_bm25_top100": ms, "ms_fuse": ms_f,
                      "vector_tflops": 2.0 * args.queries * args.docs * args.dim / (ms_v * 1e-3) / 1e12,
                      "planted_top1_ok": planted_ok})


if __name__ == "__main__":
    main()

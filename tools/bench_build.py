#!/usr/bin/env python
"""BM25 index build from tokens: the library's build kernels against a torch device-sort restatement
(what Bm25Index.postings_from_tokens did in the first version) on C3-shaped documents.

    python tools/bench_build.py --docs 2000000
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402
from super_rag_b200.bm25 import Bm25Index  # noqa: E402

VOCAB = 50_000


def torch_sort_build(tokens, doc_ptr, vocab):
    dev = tokens.device
    n_docs = doc_ptr.numel() - 1
    dl = (doc_ptr[1:] - doc_ptr[:-1]).to(torch.int32)
    doc_of_tok = torch.repeat_interleave(torch.arange(n_docs, device=dev, dtype=torch.int64), dl.to(torch.int64))
    key, _ = torch.sort(tokens.to(torch.int64) * n_docs + doc_of_tok)
    uniq, counts = torch.unique_consecutive(key, return_counts=True)
    term = torch.div(uniq, n_docs, rounding_mode="floor")
    post_doc = (uniq - term * n_docs).to(torch.int32)
    df = torch.bincount(term, minlength=vocab)
    term_ptr = torch.zeros(vocab + 1, dtype=torch.int64, device=dev)
    term_ptr[1:] = torch.cumsum(df, 0)
    return term_ptr, post_doc, counts.to(torch.int32), dl


def timed(fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    return out, time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=2_000_000)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    dl = torch.clamp(torch.round(400 + 40 * torch.randn(a.docs, generator=g, device=dev)), 50, 800).to(torch.int64)
    doc_ptr = torch.zeros(a.docs + 1, dtype=torch.int64, device=dev)
    doc_ptr[1:] = torch.cumsum(dl, 0)
    total = int(doc_ptr[-1].item())
    p = 1.0 / torch.arange(1, VOCAB + 1, dtype=torch.float64, device=dev)
    cdf = torch.cumsum(p / p.sum(), 0).to(torch.float32)
    tokens = torch.searchsorted(cdf, torch.rand(total, generator=g, device=dev)).clamp_(max=VOCAB - 1).to(torch.int32)
    ops.bm25_build_from_tokens(tokens[:10000], torch.arange(0, 10001, 400, device=dev), VOCAB)   # warm the library
    (tp, pd, ptf, dl_n, df, bp), t_native = timed(lambda: ops.bm25_build_from_tokens(tokens, doc_ptr, VOCAB))
    (tp2, pd2, ptf2, dl2), t_torch = timed(lambda: torch_sort_build(tokens, doc_ptr, VOCAB))
    same = bool(torch.equal(tp, tp2) and torch.equal(pd, pd2) and torch.equal(ptf, ptf2) and torch.equal(dl_n, dl2))
    _, t_index = timed(lambda: Bm25Index.from_postings(tp, pd, ptf, dl_n, block_ptr=bp))
    print(json.dumps({"docs": a.docs, "tokens": total, "postings": int(pd.numel()), "native_build_s": round(t_native, 3),
                      "torch_sort_build_s": round(t_torch, 3), "identical": same,
                      "impacts_and_bank_order_s": round(t_index, 3),
                      "native_tokens_per_s": round(total / t_native / 1e9, 2)}))


if __name__ == "__main__":
    main()

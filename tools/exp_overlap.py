#!/usr/bin/env python
"""C4's two scans side by side: K1 (vector top-100, tensor / power bound) on `a` SMs and K2 (BM25 top-100, LSU bound) on
148 - a SMs, on two streams, against the two back to back on all SMs.
    python tools/exp_overlap.py [--docs 10000000] [--splits 112,104,96,120]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_hybrid as bh  # noqa: E402
from super_rag_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=10_000_000)
    ap.add_argument("--splits", default="112,104,96,120")
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    idx = bh.build_index(args.docs, dev)
    qt = bh.make_queries(1024, 8, dev)
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    corpus = torch.empty(args.docs, 1024, dtype=torch.bfloat16, device=dev)
    for r0 in range(0, args.docs, 1 << 20):
        r1 = min(args.docs, r0 + (1 << 20))
        corpus[r0:r1] = torch.randn(r1 - r0, 1024, generator=g, device=dev).to(torch.bfloat16)
    _, cs = ops.ingest_rows(corpus, "cosine")
    q32 = torch.randn(1024, 1024, generator=g, device=dev)
    qb, qs = ops.ingest_rows(q32, "cosine")
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()

    def run(a):
        if a:
            os.environ["SRAG_TOPK_SMS"] = str(a)
            os.environ["SRAG_BM25_SMS"] = str(148 - a)
        else:
            os.environ.pop("SRAG_TOPK_SMS", None)
            os.environ.pop("SRAG_BM25_SMS", None)
        ws = torch.empty(ops.topk_workspace_bytes(args.docs, 1024, 100, 1024), dtype=torch.uint8, device=dev)

        def step():
            if a:
                cur = torch.cuda.current_stream()
                sa.wait_stream(cur)
                sb.wait_stream(cur)
                with torch.cuda.stream(sa):
                    v = ops.topk_search(corpus, cs, qb, qs, 100, workspace=ws)
                with torch.cuda.stream(sb):
                    b = idx.search_terms(qt, 100)
                cur.wait_stream(sa)
                cur.wait_stream(sb)
            else:
                v = ops.topk_search(corpus, cs, qb, qs, 100, workspace=ws)
                b = idx.search_terms(qt, 100)
            return v, b

        for _ in range(3):
            out = step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            out = step()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps, out

    ms0, ref = run(0)
    print(f"back to back on 148 SMs: {ms0:.2f} ms", flush=True)
    for a in [int(x) for x in args.splits.split(",")]:
        ms, out = run(a)
        same = all(torch.equal(x, y) for x, y in zip(ref[0] + ref[1], out[0] + out[1]))
        print(f"K1 on {a} SMs | K2 on {148 - a} SMs: {ms:.2f} ms  results equal: {same}", flush=True)
    ms1, _ = run(0)
    print(f"back to back again: {ms1:.2f} ms")


if __name__ == "__main__":
    main()

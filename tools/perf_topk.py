#!/usr/bin/env python
"""Quick device-side timing of K1 (not the bench contract; see bench.py for that).

    python tools/perf_topk.py --rows 10000000 --dim 1024 --batches 1024,256,64,8 --k 10
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402


def make_corpus(n, d, seed, dev, chunk=1 << 20):
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    out = torch.empty(n, d, dtype=torch.bfloat16, device=dev)
    for r0 in range(0, n, chunk):
        r1 = min(n, r0 + chunk)
        out[r0:r1] = torch.randn(r1 - r0, d, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=1024)
    ap.add_argument("--batches", default="1024")
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--group", default="")
    args = ap.parse_args()
    if args.group:
        os.environ["SRAG_TOPK_CTA_GROUP"] = args.group
    dev = torch.device("cuda:0")
    corpus = make_corpus(args.rows, args.dim, 1, dev)
    _, cs = ops.ingest_rows(corpus, "cosine")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    for B in [int(x) for x in args.batches.split(",")]:
        q = make_corpus(B, args.dim, 2, dev)
        _, qs = ops.ingest_rows(q, "cosine")
        ws = torch.empty(ops.topk_workspace_bytes(args.rows, min(B, ops.MAX_QUERIES_PER_CALL), args.k, args.dim),
                         dtype=torch.uint8, device=dev)
        for _ in range(2):
            s, i = ops.topk_search(corpus, cs, q, qs, args.k, workspace=ws)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            s, i = ops.topk_search(corpus, cs, q, qs, args.k, workspace=ws)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        flops = 2.0 * B * args.rows * args.dim
        byts = 2.0 * args.rows * args.dim
        print(json.dumps({"B": B, "k": args.k, "ms": round(ms, 3), "qps": round(B / ms * 1e3, 1),
                          "tflops": round(flops / ms / 1e9, 1), "hbm_gbs": round(byts / ms / 1e6, 1),
                          "tensor_frac": round(flops / ms / 1e9 / peaks.get("bf16_tflops", 1670.8), 3),
                          "hbm_frac": round(byts / ms / 1e6 / peaks.get("hbm_gbs", 6548.5), 3)}), flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""K1 time against shard size at B = 1024, k = 10: the intercept is the fixed cost that limits strong scaling at N = 8
(1.25 M rows per GPU). Prints total ms per search (CUDA events around the call) and the main-kernel / finalize split from
the library's own events."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402
from tools.perf_topk import make_corpus  # noqa: E402

dev = torch.device("cuda:0")
corpus = make_corpus(2_500_000, 1024, 1, dev)
_, cs = ops.ingest_rows(corpus, "cosine")
B, k = 1024, int(os.environ.get("K", "10"))
q = make_corpus(B, 1024, 2, dev)
_, qs = ops.ingest_rows(q, "cosine")
ws = torch.empty(ops.topk_workspace_bytes(2_500_000, B, k, 1024), dtype=torch.uint8, device=dev)
for _ in range(30):
    ops.topk_search(corpus, cs, q, qs, k, workspace=ws)
for n in (18944, 37888, 75776, 151552, 303104, 606208, 1_250_000, 2_500_000):
    c, s = corpus[:n], cs[:n]
    for _ in range(5):
        ops.topk_search(c, s, q, qs, k, workspace=ws)
    torch.cuda.synchronize()
    ops.prof_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    it = 20
    for _ in range(it):
        ops.topk_search(c, s, q, qs, k, workspace=ws)
    e1.record()
    torch.cuda.synchronize()
    main_ms, fin_ms, n_prof = ops.prof_read()
    ops.prof_enable(False)
    print(f"rows {n:8d} tiles/pair {n / 256 / 74:6.2f}  call {e0.elapsed_time(e1) / it:7.4f} ms  main {main_ms / n_prof:7.4f}  "
          f"finalize {fin_ms / n_prof:6.4f}", flush=True)

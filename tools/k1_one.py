#!/usr/bin/env python
"""One K1 configuration, a few launches (for ncu captures): ROWS, B, K from the environment."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402
from tools.perf_topk import make_corpus  # noqa: E402

n, B, k = int(os.environ.get("ROWS", "1250000")), int(os.environ.get("B", "1024")), int(os.environ.get("K", "100"))
dev = torch.device("cuda:0")
corpus = make_corpus(n, 1024, 1, dev)
_, cs = ops.ingest_rows(corpus, "cosine")
q = make_corpus(B, 1024, 2, dev)
_, qs = ops.ingest_rows(q, "cosine")
ws = torch.empty(ops.topk_workspace_bytes(n, B, k, 1024), dtype=torch.uint8, device=dev)
for _ in range(6):
    ops.topk_search(corpus, cs, q, qs, k, workspace=ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ops.topk_search(corpus, cs, q, qs, k, workspace=ws)
e1.record()
torch.cuda.synchronize()
print(f"rows {n} B {B} k {k}: {e0.elapsed_time(e1) / 5:.3f} ms per search")

#!/usr/bin/env python
"""A selective metadata filter on a small collection (100 k rows, the eligible rows all in ONE region so that most lists
have nothing to publish): the prescan's wait for "every list reported" must not run into its time bound."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402
from tools.perf_topk import make_corpus  # noqa: E402

dev = torch.device("cuda:0")
n, B, k = 100_000, 1000, 5
corpus = make_corpus(n, 1024, 1, dev)
_, cs = ops.ingest_rows(corpus, "cosine")
q = make_corpus(B, 1024, 2, dev)
_, qs = ops.ingest_rows(q, "cosine")
group = torch.zeros(n, dtype=torch.int32, device=dev)
group[60_000:60_500] = 7                       # 500 eligible rows in one place
qg = torch.full((B,), 7, dtype=torch.int32, device=dev)
ws = torch.empty(ops.topk_workspace_bytes(n, B, k, 1024), dtype=torch.uint8, device=dev)


def timed(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20, out


ms_u, _ = timed(lambda: ops.topk_search(corpus, cs, q, qs, k, workspace=ws))
ms_f, (s, i) = timed(lambda: ops.topk_search(corpus, cs, q, qs, k, workspace=ws, row_group=group, q_group=qg))
ok = bool(((i >= 60_000) & (i < 60_500)).all())
print(f"unfiltered {ms_u:.3f} ms   filtered (0.5 % eligible, one region) {ms_f:.3f} ms   ids inside the group: {ok}")

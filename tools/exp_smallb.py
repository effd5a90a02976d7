import sys, os, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
from super_rag_b200 import ops
import perf_topk as P
dev = torch.device("cuda:0")
corpus = P.make_corpus(10_000_000, 1024, 1, dev)
_, cs = ops.ingest_rows(corpus, "cosine")
def t(q, qs, label):
    ws = torch.empty(ops.topk_workspace_bytes(10_000_000, q.shape[0], 10, 1024), dtype=torch.uint8, device=dev)
    for _ in range(2): ops.topk_search(corpus, cs, q, qs, 10, workspace=ws)
    torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): ops.topk_search(corpus, cs, q, qs, 10, workspace=ws)
    e1.record(); torch.cuda.synchronize(); print(label, round(e0.elapsed_time(e1)/5, 3), "ms", flush=True)
for B in (1, 8, 32, 64, 96, 128):
    q = P.make_corpus(B, 1024, 2, dev); _, qs = ops.ingest_rows(q, "cosine"); t(q, qs, f"B={B} real rows")
q1 = P.make_corpus(1, 1024, 2, dev)
for pad in (8, 64, 128):
    qp = torch.zeros(pad, 1024, dtype=torch.bfloat16, device=dev); qp[:1] = q1
    _, qs = ops.ingest_rows(qp, "cosine"); t(qp, qs, f"1 real row padded with zero rows to {pad}")

import os, sys, time, torch
sys.path.insert(0, '.')
from super_rag_b200 import ops
torch.manual_seed(0)
n, d = 2_000_000, 1024
c = torch.randn(n, d, device='cuda'); cb, cs = ops.ingest_rows(c, 'cosine'); del c
for b, m in ((1024, 256), (1024, 100), (1024, 128), (256, 64)):
    q = torch.randn(b, d, device='cuda'); qb, qs = ops.ingest_rows(q, 'cosine')
    cand = torch.randint(0, n, (b, m), device='cuda')
    for _ in range(3): r = ops.mmr_rerank(cb, cs, qb, qs, cand, 0.7)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10): r = ops.mmr_rerank(cb, cs, qb, qs, cand, 0.7)
    e1.record(); torch.cuda.synchronize()
    print(os.environ.get('SRAG_MMR_TC', '1'), b, m, round(e0.elapsed_time(e1) / 10, 3), 'ms', flush=True)
    torch.save([x.cpu() for x in r], f"/tmp/mmr_{os.environ.get('SRAG_MMR_TC','1')}_{b}_{m}.pt")

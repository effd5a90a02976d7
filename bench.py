#!/usr/bin/env python
"""Headline benchmark: queries/sec at top-10 over a 10M x 1024 bf16 corpus (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch of 1024 synthetic queries: the fused tcgen05 GEMM + top-k kernel
over this rank's corpus shard, the finalize kernel (which writes the candidates as 8-byte exchange keys into the
all-gather send buffer) and, for N > 1, one NCCL all-gather + the merge kernel. For N > 1 the SAME 10M-row corpus is
row-sharded across the ranks (strong scaling; the metric is quoted on that corpus).

  value     whole-job queries/sec, inputs resident in HBM, CUDA events, max over ranks
  e2e       same metric through the public API with pinned HOST queries: H2D copy + cast/normalise + search + D2H of
            (scores, ids) every step. Headline: ShardedVectorIndex.submit_batch(...).wait() with two batches in flight
            (the copies of step i + 1 overlap the scan of step i); `one_batch_at_a_time` = search_batch + synchronize
  roofline  dominant kernel (topk_gemm_kernel) timed live with CUDA events on its stream
  cpu_baseline / --impl reference
            the CPU path timed on the host cores (oracle/cpu_baseline.py): torch (MKL) fp32 Q @ C^T blocked over rows +
            torch.topk + k-way merge, every host thread (set explicitly: torchrun exports OMP_NUM_THREADS=1), on the SAME
            seeded rows the GPU arm scans. --impl reference: K timed steps over the first 1M rows (ms_per_step is that
            measured time) plus ONE un-extrapolated pass over all 10M rows, which is what `value` reports.
  other_configs  (after the timed region; not part of `value`)
            N = 1: BASELINE.json configs[2] / [3] (BM25 top-100 over a 10M-document Zipfian index with its own HBM roofline,
            its stop-word variant, the hybrid vector + BM25 -> RRF pipeline), each with e2e, a CPU baseline and a FULL-SIZE
            oracle check (oracle/bm25.py, oracle.exact_topk_blocks, oracle.rrf_fuse_topk on >= 16 queries), and configs[0]
            (C1: 100k x 1024 fp32 through the connector API).
            N >= 2: configs[4] (C5: 100M x 1024 bf16 row-sharded over the N GPUs).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DIM = 1024
GEN_CHUNK = 65536  # rows per seeded generation chunk: corpus content is independent of the sharding
METRIC = "queries/sec at top-10, 10Mx1024 corpus"
METRIC_C5 = "queries/sec at top-10, 100Mx1024 corpus"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=1024)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--cpu-sample-rows", type=int, default=0, help="rows of the CPU sample steps (0 = 1M)")
    ap.add_argument("--e2e-depth", type=int, default=2, help="batches in flight in the e2e loop (submit_batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-full-pass", action="store_true", help="--impl reference: skip the un-extrapolated 10M-row pass")
    ap.add_argument("--no-hybrid", action="store_true", help="skip the C3 (BM25) / C4 (hybrid) lines attached at N=1")
    ap.add_argument("--no-c1", action="store_true", help="skip the C1 line attached at N=1")
    ap.add_argument("--no-c5", action="store_true", help="skip the C5 (100M rows) line attached at N>=2")
    ap.add_argument("--c5-rows", type=int, default=100_000_000)
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"], "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback"}


# ---------------------------------------------------------------------------------------------
# synthetic data: corpus N(0,1) -> bf16 generated in seeded chunks of 65536 rows (so the content does
# not depend on the sharding); queries: half fresh N(0,1), half corpus rows (from chunk 0) + 0.1 N(0,1)
# ---------------------------------------------------------------------------------------------
def gen_chunk(torch, c, dev):
    g = torch.Generator(device=dev)
    g.manual_seed(1_000_003 + c)
    return torch.randn(GEN_CHUNK, DIM, generator=g, device=dev, dtype=torch.float32)


def make_shard(torch, lo, hi, dev):
    out = torch.empty(hi - lo, DIM, dtype=torch.bfloat16, device=dev)
    for c in range(lo // GEN_CHUNK, (hi - 1) // GEN_CHUNK + 1):
        a, b = max(lo, c * GEN_CHUNK), min(hi, (c + 1) * GEN_CHUNK)
        blk = gen_chunk(torch, c, dev)
        out[a - lo:b - lo] = blk[a - c * GEN_CHUNK:b - c * GEN_CHUNK].to(torch.bfloat16)
    return out


def make_queries(torch, n_q, n_rows, dev):
    g = torch.Generator(device=dev)
    g.manual_seed(2)
    q = torch.randn(n_q, DIM, generator=g, device=dev, dtype=torch.float32)
    h = n_q // 2
    if h and n_rows >= 1:
        src = torch.randint(0, min(GEN_CHUNK, n_rows), (h,), generator=g, device=dev)
        q[:h] = gen_chunk(torch, 0, dev)[src] + 0.1 * q[:h]
        return q, src.cpu()
    return q, None


class ClockSampler:
    """nvidia-smi sampled DURING the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index):
        self.proc = None
        self.index = index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# CPU arm (oracle/cpu_baseline.py; the oracle package is only ever the checker or the timed CPU baseline)
# ---------------------------------------------------------------------------------------------
def cpu_blocks(torch, rows, gen_dev):
    """The first `rows` rows of the benchmark corpus as fp32 host blocks holding the bf16-rounded values the GPU arm
    scans: generated by the very generator of the GPU arm when a CUDA device is present (so both arms see the same
    data), by torch's CPU generator otherwise (same distribution, other numbers; the line says which)."""
    for c in range((rows + GEN_CHUNK - 1) // GEN_CHUNK):
        blk = gen_chunk(torch, c, gen_dev)[:max(0, min(GEN_CHUNK, rows - c * GEN_CHUNK))]
        yield c * GEN_CHUNK, blk.to(torch.bfloat16).to(torch.float32).cpu()


def time_cpu_arm(args, sample_rows, steps, warmup, full_pass):
    """-> (cpu_baseline dict, measured seconds of one sample step, full-pass dict or None)."""
    import torch
    from oracle import cpu_baseline as cb
    threads = cb.use_all_threads()
    gen_dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if torch.cuda.is_available() else torch.device("cpu")
    same = gen_dev.type == "cuda"
    q32, _ = make_queries(torch, args.queries, args.rows, gen_dev)
    q = q32.to(torch.bfloat16).to(torch.float32).cpu()
    q_scale = cb.inv_norms(q)
    sample_rows = min(sample_rows, args.rows)
    blocks, scales, row0 = [], [], []
    for r0, blk in cpu_blocks(torch, sample_rows, gen_dev):
        blocks.append(blk); scales.append(cb.inv_norms(blk)); row0.append(r0)
    t_step, _ = cb.time_topk(blocks, scales, row0, q, q_scale, args.k, steps, warmup)
    frac = sample_rows / args.rows
    value = args.queries * frac / t_step                   # cost is linear in the rows scanned
    full = None
    if full_pass and sample_rows < args.rows:
        del blocks, scales
        t_full, rows_done, _ = cb.time_topk_streamed(cpu_blocks(torch, args.rows, gen_dev), q, q_scale, args.k)
        full = {"rows": rows_done, "seconds": t_full, "queries_per_s": args.queries / t_full, "extrapolated": False,
                "sample_rate_agrees_within": abs(args.queries / t_full - value) / value}
        value = args.queries / t_full
    base = {"value": value, "unit": "queries/s", "cores": threads, "threads": threads, "cpu_model": cb.cpu_model(),
            "host_cpus": os.cpu_count(), "kind": "port", "extrapolated": full is None and sample_rows < args.rows,
            "same_data_as_gpu_arm": same,
            "sample": f"{args.queries} queries x first {sample_rows} rows of the benchmark corpus "
                      f"({'same seeded chunks as the GPU arm' if same else 'CPU generator: no CUDA device here'}), "
                      f"torch fp32 matmul (MKL, {threads} threads) + torch.topk per 65536-row block + k-way merge, "
                      f"{steps} steps of {t_step:.3f} s"
                      + (f"; value = ONE un-extrapolated pass over all {args.rows} rows in {full['seconds']:.1f} s"
                         if full else f"; value scaled x{1 / frac:.1f} to {args.rows} rows")}
    return base, t_step, full


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_rows = args.cpu_sample_rows or 1_048_576
    base, t_step, full = time_cpu_arm(args, sample_rows, args.steps, max(1, args.warmup), not args.no_full_pass)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"],
            "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": t_step * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic",
            "config": {"workload": f"C2: {args.rows}x{DIM} corpus (bf16-rounded values), {args.queries} queries, "
                                   f"top-{args.k}, exact cosine on the host cores",
                       "step": f"one pass over the first {min(sample_rows, args.rows)} rows (ms_per_step is that measured time)",
                       "full_pass": full},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# hooks for tools/bench_hybrid.py: full-size oracle checks and CPU baselines for C3 / C4
# ---------------------------------------------------------------------------------------------
def _mini_index(ctx, q_rows):
    """Host CSR over just the terms of the check queries, impacts recomputed by oracle/bm25.py from (tf, dl) with the GLOBAL
    statistics of the full index. Term ids are remapped in ascending order, so the summation order is unchanged."""
    import numpy as np
    import oracle
    idx, kept = ctx["idx"], ctx["kept"]
    terms = sorted({int(t) for r in q_rows for t in r if t >= 0})
    remap = {t: j for j, t in enumerate(terms)}
    docs = [kept["terms"][t][0] for t in terms]
    tfs = [kept["terms"][t][1] for t in terms]
    tp = np.zeros(len(terms) + 1, dtype=np.int64)
    np.cumsum([len(d) for d in docs], out=tp[1:])
    pd = np.concatenate(docs) if docs else np.zeros(0, np.int32)
    ptf = np.concatenate(tfs) if tfs else np.zeros(0, np.int32)
    df_sel = np.asarray(idx.stats.df)[terms]
    imp = oracle.bm25_impacts(tp, pd, ptf, kept["dl"], n_docs_global=idx.stats.n_docs, df_global=df_sel, avgdl=idx.stats.avgdl)
    queries = [[remap[int(t)] for t in r if t >= 0] for r in q_rows]
    return tp, pd, imp, queries


def hook_c3_check(ctx):
    """C3 at its full size against oracle/bm25.py: impacts recomputed on the host, ids and score BITS compared."""
    import numpy as np
    import oracle
    n = ctx["n_check"]
    q_rows = ctx["qt"][:n].cpu().numpy().tolist()
    tp, pd, imp, queries = _mini_index(ctx, q_rows)
    want_s, want_i = oracle.bm25_topk(tp, pd, imp, ctx["n_docs"], queries, ctx["k"])
    got_s, got_i = ctx["scores"][:n].cpu().numpy(), ctx["ids"][:n].cpu().numpy()
    ids_ok = bool((got_i == want_i).all())
    bits_ok = bool((got_s.view(np.uint32) == want_s.view(np.uint32)).all())
    ctx["_mini"] = (tp, pd, imp, queries, want_i)
    return {"queries": n, "ok": ids_ok and bits_ok, "ids_equal": ids_ok, "score_bits_equal": bits_ok,
            "checker": "oracle/bm25.py (impacts recomputed from tf, dl and the global statistics on the host)"}


def hook_c3_cpu(ctx):
    from oracle import cpu_baseline as cb
    if "_mini" not in ctx:
        hook_c3_check(ctx)
    tp, pd, imp, queries, _ = ctx["_mini"]
    t_q = cb.time_bm25(tp, pd, imp, ctx["n_docs"], queries, ctx["k"])
    return {"value": 1.0 / t_q, "unit": "queries/s", "cores": 1, "kind": "port", "cpu_model": cb.cpu_model(),
            "sample": f"{len(queries)} of the batch's queries: numpy CSR gather + accumulate + partition top-{ctx['k']} "
                      f"(oracle.bm25_topk), {t_q * 1e3:.1f} ms per query, one core"}


def _device_blocks(torch, corpus, block_rows=262144):
    """Stream a device bf16 corpus to the host as fp32 blocks through a pinned staging buffer."""
    stage = torch.empty(block_rows, corpus.shape[1], dtype=torch.bfloat16).pin_memory()
    for r0 in range(0, corpus.shape[0], block_rows):
        m = min(block_rows, corpus.shape[0] - r0)
        stage[:m].copy_(corpus[r0:r0 + m], non_blocking=False)
        yield r0, stage[:m].to(torch.float32).numpy()


def hook_c4_check(ctx):
    """C4 at its full size: the oracle's vector top-100 (fp32 BLAS over the whole 10M-row corpus streamed to the host) and
    BM25 top-100 of the check queries, fused by oracle.rrf_fuse_topk, against the device pipeline's top-10."""
    import numpy as np
    import torch
    import oracle
    n = ctx["n_check"]
    if "_mini" not in ctx:
        hook_c3_check(ctx)
    want_bm25 = ctx["_mini"][4]
    qb = ctx["q32"][:n].to(torch.bfloat16).to(torch.float32).cpu().numpy()
    vs, vi = oracle.exact_topk_blocks(_device_blocks(torch, ctx["corpus"]), qb, 100, "cosine")
    got_vi, got_vs = ctx["vec_ids"][:n].cpu().numpy(), ctx["vec_scores"][:n].cpu().numpy()
    vec_same = int((got_vi == vi).all(axis=1).sum())
    score_err = float(np.abs(got_vs - vs).max())
    # ids may only differ where the oracle's own neighbouring scores are closer than the north-star tolerance
    explained = True
    for q in range(n):
        for j in np.nonzero(got_vi[q] != vi[q])[0]:
            lo, hi = max(0, j - 1), min(99, j + 1)
            explained &= bool(min(abs(vs[q, j] - vs[q, lo]) if lo != j else 1.0, abs(vs[q, j] - vs[q, hi]) if hi != j else 1.0) <= 1e-3)
    fused_ok, fused_ok_given_lists = 0, 0
    got_fi, got_fs = ctx["fused_ids"][:n].cpu().numpy(), ctx["fused_scores"][:n].cpu().numpy()
    for q in range(n):
        wi, ws = oracle.rrf_fuse_topk([vi[q], want_bm25[q]], ctx["rank_const"], 0.0, ctx["k_out"])
        fused_ok += int(got_fi[q, :len(wi)].tolist() == wi.tolist() and got_fs[q, :len(ws)].tolist() == ws.tolist())
        wi2, ws2 = oracle.rrf_fuse_topk([got_vi[q], want_bm25[q]], ctx["rank_const"], 0.0, ctx["k_out"])
        fused_ok_given_lists += int(got_fi[q, :len(wi2)].tolist() == wi2.tolist() and got_fs[q, :len(ws2)].tolist() == ws2.tolist())
    ctx["_c4_lists"] = [(vi[q].tolist(), want_bm25[q].tolist()) for q in range(n)]
    return {"queries": n, "ok": bool(fused_ok_given_lists == n and explained and score_err <= 1e-3),
            "vector_top100_identical_to_oracle": vec_same, "vector_id_differences_within_tolerance": explained,
            "vector_max_abs_score_error": score_err, "fused_top10_equal_oracle_end_to_end": fused_ok,
            "fused_top10_equal_oracle_rrf_of_device_lists": fused_ok_given_lists,
            "checker": "oracle.exact_topk_blocks over all rows + oracle/bm25.py + oracle.rrf_fuse_topk"}


def hook_c4_cpu(ctx):
    """CPU arm of the hybrid pipeline = CPU vector top-100 (torch arm, first 1M rows, scaled) + CPU BM25 + Python rrf."""
    import torch
    from oracle import cpu_baseline as cb
    threads = cb.use_all_threads()
    n_rows = ctx["corpus"].shape[0]
    sample = min(1 << 20, n_rows)
    blk = ctx["corpus"][:sample].to(torch.float32).cpu()
    blocks = list(torch.split(blk, GEN_CHUNK))
    scales = [cb.inv_norms(b) for b in blocks]
    row0 = list(range(0, sample, GEN_CHUNK))
    q = ctx["q32"].to(torch.bfloat16).to(torch.float32).cpu()
    t_vec, _ = cb.time_topk(blocks, scales, row0, q, cb.inv_norms(q), 100, steps=1, warmup=1)
    t_vec_q = t_vec * (n_rows / sample) / q.shape[0]
    tp, pd, imp, queries, _ = ctx["_mini"]
    t_bm = cb.time_bm25(tp, pd, imp, ctx["n_docs"], queries, 100)
    t_rrf = cb.time_rrf(ctx.get("_c4_lists") or [([1, 2, 3], [3, 4, 5])], ctx["rank_const"], ctx["k_out"])
    t_q = t_vec_q + t_bm + t_rrf
    return {"value": 1.0 / t_q, "unit": "queries/s", "cores": threads, "kind": "port", "cpu_model": cb.cpu_model(),
            "ms_per_query": {"vector_top100": t_vec_q * 1e3, "bm25_top100": t_bm * 1e3, "rrf": t_rrf * 1e3},
            "sample": f"vector: torch matmul + topk over the first {sample} rows x {q.shape[0]} queries ({threads} threads), scaled "
                      f"x{n_rows / sample:.1f}; BM25: {len(queries)} queries numpy CSR, one core; rrf: the reference's Python "
                      f"function restated (search_utils.py:1763-1778), {len(ctx.get('_c4_lists') or [])} queries"}


def run_c1(torch, dev):
    """BASELINE.json configs[0]: exact cosine top-5, 100k x 1024 fp32, 1k queries, through the connector search API —
    GPU connector (single-query calls and one batched call) next to the CPU paths (tools/bench_c1.py)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_c1
    return bench_c1.run(dev)


def run_c5(torch, dist, ops, ShardedVectorIndex, shard_range, args, world, rank, dev, barrier, max_over_ranks):
    """BASELINE.json configs[4]: 100M x 1024 bf16 (204.8 GB) row-sharded over the N GPUs, same step as the headline."""
    rows = args.c5_rows
    lo, hi = shard_range(rows, world, rank)
    free, _ = torch.cuda.mem_get_info(dev)
    need = (hi - lo) * (DIM * 2 + 4) + (4 << 30)
    fits = torch.tensor([int(free > need)], device=dev)
    dist.all_reduce(fits, op=dist.ReduceOp.MIN)
    if not bool(fits.item()):
        return {"config": "C5", "error": f"{rows} rows need {need / 1e9:.0f} GB per GPU at {world} GPUs, {free / 1e9:.0f} GB free"}
    corpus = make_shard(torch, lo, hi, dev)
    _, row_scale = ops.ingest_rows(corpus, "cosine")
    index = ShardedVectorIndex(corpus, row_scale, id_base=lo, metric="cosine")
    q32, planted = make_queries(torch, args.queries, rows, dev)
    qb, qs = ops.ingest_rows(q32, "cosine")
    B, k = args.queries, args.k
    steps = max(3, min(args.steps, 5))
    for _ in range(2):
        s, i = index.search_device(qb, qs, k)
    ops.prof_enable(True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        s, i = index.search_device(qb, qs, k)
    e1.record()
    barrier()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / steps
    main_ms, fin_ms, n_prof = ops.prof_read()
    ops.prof_enable(False)
    if n_prof == 0:   # the step ran as a CUDA graph replay: time K1 on a few eager steps instead
        ops.prof_enable(True)
        for _ in range(3):
            index._step(qb, qs, k)
        barrier()
        main_ms, fin_ms, n_prof = ops.prof_read()
        ops.prof_enable(False)
    k1_ms = max_over_ranks(main_ms / max(1, n_prof))
    ok = bool((i[:planted.numel(), 0].cpu() == planted).all()) if planted is not None else None
    q_host = q32.cpu().pin_memory()
    out_s = torch.empty(B, k, dtype=torch.float32).pin_memory()
    out_i = torch.empty(B, k, dtype=torch.int64).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        s2, i2 = index.search_batch(q_host, k)
        out_s.copy_(s2, non_blocking=True); out_i.copy_(i2, non_blocking=True)
        torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0) / steps
    pk = peaks()
    n_local = hi - lo
    tfl = 2.0 * B * n_local * DIM / (k1_ms * 1e-3) / 1e12
    index.close()
    del index, corpus, row_scale
    torch.cuda.empty_cache()
    return {"config": "C5", "metric": METRIC_C5, "value": B / (ms_step * 1e-3), "unit": "queries/s", "n_gpus": world,
            "steps": steps, "ms_per_step": ms_step, "scaling": "strong",
            "workload": f"C5: {rows}x{DIM} bf16 corpus ({rows * DIM * 2 / 1e9:.1f} GB) row-sharded over {world} GPUs, {B} queries/step, "
                        f"top-{k}, exact cosine; per-shard top-k -> 8-byte exchange keys -> NCCL all-gather -> merge",
            "rows_per_gpu": n_local, "shard_gb": n_local * DIM * 2 / 1e9,
            "roofline": {"bound": "tensor", "kernel": "topk_gemm_kernel", "kernel_ms": k1_ms, "achieved": tfl,
                         "peak": pk["bf16_tflops"], "unit": "TFLOP/s per GPU", "frac": tfl / pk["bf16_tflops"]},
            "exchange_and_merge_us": (ms_step - k1_ms) * 1e3,
            "e2e": {"value": B / e2e_s, "unit": "queries/s", "ms_per_step": e2e_s * 1e3, "h2d_bytes_per_step": B * DIM * 4,
                    "d2h_bytes_per_step": B * k * 12, "api": "search_batch + D2H + synchronize, one batch at a time"},
            "planted_top1_ok": ok}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from super_rag_b200 import ops
    from super_rag_b200.sharded import ShardedVectorIndex, shard_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus and rank == 0:
        print(f"# note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lo, hi = shard_range(args.rows, world, rank)
    corpus = make_shard(torch, lo, hi, dev)
    _, row_scale = ops.ingest_rows(corpus, "cosine")
    index = ShardedVectorIndex(corpus, row_scale, id_base=lo, metric="cosine")
    q32, planted = make_queries(torch, args.queries, args.rows, dev)
    qb, qs = ops.ingest_rows(q32, "cosine")
    q_host = q32.cpu().pin_memory()
    B, k = args.queries, args.k

    # ---- device-resident timing ---------------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        s, i = index.search_device(qb, qs, k)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    ops.prof_enable(True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        s, i = index.search_device(qb, qs, k)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    main_ms, fin_ms, n_prof = ops.prof_read()
    ops.prof_enable(False)
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = B / (ms_step * 1e-3)
    graph_step = bool(getattr(index, "use_graph", False) and index._graphs)
    exchange_kind = None
    if world > 1:
        exchange_kind = ("8-byte keys from the finalize kernel -> ONE kernel: P2P stores into every rank's symmetric buffer over "
                         "NVLink + flags + merge (no collective call)" if getattr(index, "p2p", None) is not None else
                         "8-byte keys from the finalize kernel -> NCCL all_gather -> merge_keys" if index.use_keys else
                         "two int64 planes -> NCCL all_gather -> merge_candidates")
    if n_prof == 0:
        # the step ran as a CUDA graph replay (the library's timing events are not part of the graph): time the dominant
        # kernel on a few eager steps right after the timed region instead
        ops.prof_enable(True)
        for _ in range(5):
            index._step(qb, qs, k)
        barrier()
        main_ms, fin_ms, n_prof = ops.prof_read()
        ops.prof_enable(False)

    # the dominant kernel's time on every rank (the step waits for the slowest at the exchange)
    k1_per_rank = [main_ms / max(1, n_prof)]
    if world > 1:
        t = torch.tensor([main_ms / max(1, n_prof)], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        k1_per_rank = [float(x.item()) for x in allt]

    # sanity: planted queries must retrieve their source row first (cheap, no oracle)
    ok = True
    if planted is not None:
        top1 = i[:planted.numel(), 0].cpu()
        ok = bool((top1 == planted).all())

    # ---- end to end through the public API, host buffers ---------------------------------------
    out_s = torch.empty(B, k, dtype=torch.float32).pin_memory()
    out_i = torch.empty(B, k, dtype=torch.int64).pin_memory()
    for _ in range(max(1, args.warmup)):
        s2, i2 = index.search_batch(q_host, k)
        out_s.copy_(s2, non_blocking=True); out_i.copy_(i2, non_blocking=True)
    barrier()
    # (a) one batch at a time: submit, copy back, wait — the GPU idles while the host prepares the next call
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s2, i2 = index.search_batch(q_host, k)
        out_s.copy_(s2, non_blocking=True); out_i.copy_(i2, non_blocking=True)
        torch.cuda.synchronize()
    e2e_serial_s = max_over_ranks(time.perf_counter() - t0) / args.steps
    barrier()
    # (b) the same calls with two batches in flight (submit i + 1, then wait for i): every step still copies its 4 MiB
    # of fp32 queries from pinned host memory and its results back, inside the timed region; the copies of step i + 1
    # overlap the scan of step i. This is the e2e figure; (a) is reported beside it.
    depth = max(1, args.e2e_depth)
    outs = [(out_s, out_i)] + [(torch.empty_like(out_s).pin_memory(), torch.empty_like(out_i).pin_memory())
                               for _ in range(depth - 1)]
    index.submit_batch(q_host, k, *outs[0]).wait()
    barrier()
    t0 = time.perf_counter()
    pend = []
    for it in range(args.steps):
        pend.append(index.submit_batch(q_host, k, *outs[it % depth]))
        if len(pend) >= depth:
            pend.pop(0).wait()
    for p_ in pend:
        p_.wait()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0) / args.steps
    e2e_ok = bool(all(torch.equal(o[1], i.cpu()) for o in outs[:min(depth, args.steps)]))
    barrier()

    line = None
    if rank == 0:
        pk = peaks()
        n_local = hi - lo
        t_main = main_ms / max(1, n_prof) * 1e-3
        flops = 2.0 * B * n_local * DIM
        byts = 2.0 * n_local * DIM + 2.0 * B * DIM + 12.0 * B * k
        ridge = pk["bf16_tflops"] * 1e12 / (pk["hbm_gbs"] * 1e9)   # flop per corpus byte
        if B >= ridge:
            roof = {"bound": "tensor", "achieved": flops / t_main / 1e12, "peak": pk["bf16_tflops"], "unit": "TFLOP/s"}
        else:
            roof = {"bound": "hbm", "achieved": byts / t_main / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s"}
        roof["frac"] = roof["achieved"] / roof["peak"]
        roof["peak_source"] = f"{pk['source']} (burst figure: kernel timed alone with CUDA events)"
        roof["kernel"] = "topk_gemm_kernel"
        roof["kernel_ms"] = t_main * 1e3
        roof["finalize_ms"] = fin_ms / max(1, n_prof)
        roof["kernel_ms_per_rank"] = k1_per_rank
        roof["algorithmic_flop_per_launch"] = flops
        roof["algorithmic_bytes_per_launch"] = byts
        roof["traffic"] = None
        roof["traffic_source"] = None
        try:
            with open(os.path.join(ROOT, "profiles", "topk_traffic.json")) as f:
                tj = json.load(f)
            roof["traffic"] = tj.get(f"dram_bytes_per_launch_B{B}_N{n_local}")
            if roof["traffic"] is not None:
                roof["traffic_source"] = tj.get("source", "ncu --set full capture committed under profiles/ (not measured in this run)")
        except Exception:
            pass
        line = {"metric": METRIC, "value": value, "unit": "queries/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic",
                "config": {"workload": f"C2: {args.rows}x{DIM} bf16 corpus row-sharded over {world} GPU(s), "
                                       f"{B} queries/step, top-{k}, exact cosine (fp32 accumulate)",
                           "rows_per_gpu": n_local, "l2": "inputs larger than L2 (corpus shard "
                           f"{2.0 * n_local * DIM / 1e9:.2f} GB vs 126 MB L2), no flush needed",
                           "planted_top1_ok": ok, "exchange": exchange_kind, "cuda_graph_step": graph_step},
                "roofline": roof,
                "e2e": {"value": B / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": B * DIM * 4,
                        "d2h_bytes_per_step": B * k * 12, "ms_per_step": e2e_s * 1e3,
                        "api": "ShardedVectorIndex.submit_batch(pinned host queries, k, pinned host outputs).wait(), "
                               f"{depth} batches in flight", "batches_in_flight": depth,
                        "one_batch_at_a_time": {"value": B / e2e_serial_s, "ms_per_step": e2e_serial_s * 1e3,
                                                "api": "search_batch + D2H + synchronize per step"},
                        "same_ids_as_device_run": e2e_ok},
                "gpu_launches": args.steps * (2 + (1 if world > 1 else 0)),
                "clocks": clocks}
    index.close()
    del index
    other = []
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        try:
            base, _, _ = time_cpu_arm(args, args.cpu_sample_rows or 1_048_576, steps=2, warmup=1, full_pass=False)
            line["cpu_baseline"] = base
        except Exception as e:
            line["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"}
    if world == 1 and not args.no_hybrid and args.rows == 10_000_000:
        # BASELINE.json configs[2] / configs[3] beside the headline (not part of `value`)
        try:
            del corpus, row_scale
            torch.cuda.empty_cache()
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_hybrid
            hooks = bench_hybrid.Hooks()
            hooks.c3_check, hooks.c3_cpu, hooks.c4_check, hooks.c4_cpu = hook_c3_check, hook_c3_cpu, hook_c4_check, hook_c4_cpu
            hy = argparse.Namespace(docs=args.rows, queries=B, terms=8, steps=5, warmup=3, check=16, skip_c4=False,
                                    skip_stopwords=False, dim=DIM)
            other += list(bench_hybrid.run(hy, dev, hooks))
        except Exception as e:  # the headline line must still be printed
            other.append({"config": "C3/C4", "error": f"{type(e).__name__}: {e}"})
        torch.cuda.empty_cache()
    if world == 1 and not args.no_c1:
        try:
            other.append(run_c1(torch, dev))
        except Exception as e:
            other.append({"config": "C1", "error": f"{type(e).__name__}: {e}"})
    if world > 1 and not args.no_c5:
        try:
            corpus = row_scale = None
            torch.cuda.empty_cache()
            c5 = run_c5(torch, dist, ops, ShardedVectorIndex, shard_range, args, world, rank, dev, barrier, max_over_ranks)
            other.append(c5)
        except Exception as e:
            other.append({"config": "C5", "error": f"{type(e).__name__}: {e}"})
    if rank == 0:
        if other:
            line["other_configs"] = other
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

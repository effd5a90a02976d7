#!/usr/bin/env python
"""Headline benchmark: queries/sec at top-10 over a 10M x 1024 bf16 corpus (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch of 1024 synthetic queries: the fused
tcgen05 GEMM + top-k kernel over this rank's corpus shard, the finalize kernel and, for N > 1,
the NCCL all-gather of (score, id) candidates + the merge kernel. For N > 1 the SAME 10M-row
corpus is row-sharded across the ranks (strong scaling; the metric is quoted on that corpus).

  value     whole-job queries/sec, inputs resident in HBM, CUDA events, max over ranks
  e2e       same metric through the public API (ShardedVectorIndex.search_batch) with pinned
            HOST queries: H2D copy + cast/normalise + search + D2H of (scores, ids) every step
  roofline  dominant kernel (topk_gemm_kernel) timed live with CUDA events on its stream
  cpu_baseline / --impl reference   the CPU oracle port (numpy BLAS, all host threads) on a
            bounded row sample of the same workload, extrapolated linearly in rows
  other_configs  (N = 1 only, after the timed region; not part of `value`) BASELINE.json configs[2] and [3]: BM25 top-100 over
            a 10 M-document Zipfian index with its own HBM roofline, and the hybrid vector + BM25 -> RRF pipeline
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=1024)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--cpu-sample-rows", type=int, default=0, help="rows of the CPU baseline sample (0 = default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-hybrid", action="store_true", help="skip the C3 (BM25) / C4 (hybrid) lines attached at N=1")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"], "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback"}


# ---------------------------------------------------------------------------------------------
# synthetic data: corpus N(0,1) -> bf16 generated on the device in seeded chunks;
# queries: half fresh N(0,1), half corpus rows (from chunk 0) + 0.1 N(0,1)
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
# CPU arm: the oracle port on a bounded row sample (the oracle is only ever the checker / the
# timed CPU baseline; nothing in super_rag_b200/ imports it)
# ---------------------------------------------------------------------------------------------
def cpu_sample_inputs(n_rows, n_q):
    import numpy as np
    rng = np.random.default_rng(1)
    corpus = rng.standard_normal((n_rows, DIM), dtype=np.float32)
    q = np.random.default_rng(2).standard_normal((n_q, DIM), dtype=np.float32)
    h = n_q // 2
    src = np.random.default_rng(3).integers(0, n_rows, size=h)
    q[:h] = corpus[src] + 0.1 * q[:h]
    return corpus, q


def time_cpu_oracle(args, sample_rows, steps, warmup):
    import oracle
    corpus, q = cpu_sample_inputs(sample_rows, args.queries)
    corpus = oracle.bf16_round(corpus)
    q = oracle.bf16_round(q)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.exact_topk(corpus, q, args.k, "cosine")
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    scale = args.rows / sample_rows                      # cost is linear in rows scanned
    cores = os.cpu_count() or 1
    return {"value": args.queries / (t * scale), "unit": "queries/s", "cores": cores, "kind": "port",
            "sample": f"{args.queries} queries x {sample_rows}-row slice (same generator shape), fp32 numpy BLAS + "
                      f"partition top-{args.k}, {t:.2f} s/step, extrapolated x{scale:.1f} to {args.rows} rows"}, t * scale


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_rows = args.cpu_sample_rows or 131072
    base, step_s = time_cpu_oracle(args, sample_rows, args.steps, args.warmup)
    line = {"impl": "reference", "metric": "queries/sec at top-10, 10Mx1024 corpus", "value": base["value"],
            "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic",
            "config": {"workload": f"C2: {args.rows}x{DIM} corpus (bf16-rounded values), {args.queries} queries, "
                                   f"top-{args.k}, exact cosine; CPU oracle port on a row sample"},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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
    for _ in range(args.warmup):
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
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s2, i2 = index.search_batch(q_host, k)
        out_s.copy_(s2, non_blocking=True); out_i.copy_(i2, non_blocking=True)
        torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0) / args.steps
    barrier()

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
        roof["traffic"] = None
        try:
            with open(os.path.join(ROOT, "profiles", "topk_traffic.json")) as f:
                roof["traffic"] = json.load(f).get(f"dram_bytes_per_launch_B{B}_N{n_local}")
        except Exception:
            pass
        line = {"metric": "queries/sec at top-10, 10Mx1024 corpus", "value": value, "unit": "queries/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic",
                "config": {"workload": f"C2: {args.rows}x{DIM} bf16 corpus row-sharded over {world} GPU(s), "
                                       f"{B} queries/step, top-{k}, exact cosine (fp32 accumulate)",
                           "rows_per_gpu": n_local, "l2": "inputs larger than L2 (corpus shard "
                           f"{2.0 * n_local * DIM / 1e9:.2f} GB vs 126 MB L2), no flush needed",
                           "planted_top1_ok": ok},
                "roofline": roof,
                "e2e": {"value": B / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": B * DIM * 4,
                        "d2h_bytes_per_step": B * k * 12, "ms_per_step": e2e_s * 1e3},
                "gpu_launches": args.steps * (2 + (1 if world > 1 else 0)),
                "clocks": clocks}
        if world == 1 and not args.no_cpu_baseline:
            base, _ = time_cpu_oracle(args, args.cpu_sample_rows or 262144, steps=2, warmup=1)
            line["cpu_baseline"] = base
        if world == 1 and not args.no_hybrid and args.rows == 10_000_000:
            # BASELINE.json configs[2] / configs[3] beside the headline (not part of `value`): BM25 top-100 over a 10 M-document
            # Zipfian index and the hybrid vector + BM25 -> RRF pipeline, through tools/bench_hybrid.py
            try:
                torch.cuda.empty_cache()
                sys.path.insert(0, os.path.join(ROOT, "tools"))
                import bench_hybrid
                hy = argparse.Namespace(docs=args.rows, queries=B, terms=8, steps=5, warmup=3, check=2, skip_c4=False, dim=DIM)
                line["other_configs"] = list(bench_hybrid.run(hy, dev))
            except Exception as e:  # the headline line must still be printed
                line["other_configs"] = {"error": f"{type(e).__name__}: {e}"}
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

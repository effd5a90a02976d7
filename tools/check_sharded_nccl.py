#!/usr/bin/env python
"""N-GPU parity check over NCCL (run under torchrun on a multi-GPU box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tools/check_sharded_nccl.py

Every rank builds the same seeded corpus, keeps only its row shard, and the sharded search
(K1 per shard -> NCCL all-gather -> K4 merge) must return, on every rank, bit-identical ids and scores
to rank 0's single-GPU scan of the whole corpus. Also checks BM25 doc-range shards with global
statistics against the single-GPU index.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from super_rag_b200 import ops  # noqa: E402
from super_rag_b200.bm25 import Bm25Index, Bm25Stats  # noqa: E402
from super_rag_b200.sharded import (ShardedBm25Index, ShardedVectorIndex, exchange_candidates, hybrid_search,  # noqa: E402
                                    shard_range)


def mark(rank, what):
    if rank == 0:
        print(f"[check_sharded_nccl] {what}", flush=True)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    n, d, b, k = 200_003, 1024, 300, 10
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    C = torch.randn(n, d, generator=g, device=dev)
    Q = torch.randn(b, d, generator=g, device=dev)
    Q[:b // 2] = C[torch.randint(0, n, (b // 2,), generator=g, device=dev)] + 0.1 * Q[:b // 2]
    C[150_000] = C[11]                      # a tie across shards
    Q[0] = C[11]
    cb, cs = ops.ingest_rows(C, "cosine")
    lo, hi = shard_range(n, world, rank)
    index = ShardedVectorIndex(cb[lo:hi].contiguous(), cs[lo:hi].contiguous(), id_base=lo)
    s, i = index.search_batch(Q, k)
    qb, qs = ops.ingest_rows(Q, "cosine")
    fs, fi = ops.topk_search(cb, cs, qb, qs, k)
    torch.cuda.synchronize()
    ok_v = bool(torch.equal(i, fi)) and bool(torch.equal(s.view(torch.int32), fs.view(torch.int32)))
    assert i[0, 0].item() == 11 and i[0, 1].item() == 150_000
    assert index.use_keys, "global rows fit 32 bits: the 8-byte exchange keys must be in use"
    # the same step captured into a CUDA graph (SRAG_SHARDED_GRAPH=1): replayed twice, second time with other queries
    index_g = ShardedVectorIndex(cb[lo:hi].contiguous(), cs[lo:hi].contiguous(), id_base=lo)
    index_g.use_graph = True
    gs1, gi1 = index_g.search_batch(Q, k)
    gs1, gi1 = gs1.clone(), gi1.clone()
    Q2 = Q.flip(0).contiguous()
    gs2, gi2 = index_g.search_batch(Q2, k)
    gs2, gi2 = gs2.clone(), gi2.clone()
    gs3, gi3 = index_g.search_batch(Q, k)
    torch.cuda.synchronize()
    graph_used = index_g.use_graph and len(index_g._graphs) == 1
    index_g.close()     # captured NCCL work must be gone before destroy_process_group() (it hangs otherwise)
    del index_g
    mark(rank, f"vector paths done (graph captured: {graph_used})")
    ok_g = bool(torch.equal(gi1, fi)) and bool(torch.equal(gs1.view(torch.int32), fs.view(torch.int32))) and \
        bool(torch.equal(gi2, fi.flip(0))) and bool(torch.equal(gi3, fi)) and bool(torch.equal(gs3.view(torch.int32), fs.view(torch.int32)))
    ok_v = ok_v and ok_g
    # the one-kernel exchange over peer memory (SRAG_SHARDED_P2P=1): three steps (both parities, a second batch shape)
    os.environ["SRAG_SHARDED_P2P"] = "1"
    index_p = ShardedVectorIndex(cb[lo:hi].contiguous(), cs[lo:hi].contiguous(), id_base=lo)
    os.environ.pop("SRAG_SHARDED_P2P")
    p2p_used = index_p.p2p is not None
    if p2p_used:
        for rep in range(3):
            ps, pi = index_p.search_batch(Q, k)
            ok_v = ok_v and bool(torch.equal(pi, fi)) and bool(torch.equal(ps.view(torch.int32), fs.view(torch.int32)))
        ps, pi = index_p.search_batch(Q[:77], 5)
        ws5, wi5 = ops.topk_search(cb, cs, *ops.ingest_rows(Q[:77], "cosine"), 5)
        torch.cuda.synchronize()
        ok_v = ok_v and bool(torch.equal(pi, wi5)) and bool(torch.equal(ps.view(torch.int32), ws5.view(torch.int32)))
    mark(rank, f"p2p exchange done (used: {p2p_used})")
    # ids beyond the 32-bit key range take the two-plane exchange: same answers
    big = ShardedVectorIndex(cb[lo:hi].contiguous(), cs[lo:hi].contiguous(), id_base=lo + (1 << 33))
    assert not big.use_keys
    bs2, bi2 = big.search_batch(Q, k)
    torch.cuda.synchronize()
    ok_v = ok_v and bool(torch.equal(bi2 - (1 << 33), fi)) and bool(torch.equal(bs2.view(torch.int32), fs.view(torch.int32)))

    # BM25: doc-range shards with GLOBAL statistics == single index
    n_docs, vocab = 120_000, 3000
    rng = np.random.default_rng(3)
    dl = rng.integers(20, 80, n_docs)
    p = 1.0 / np.arange(1, vocab + 1)
    tokens = rng.choice(vocab, size=int(dl.sum()), p=p / p.sum()).astype(np.int64)
    doc_ptr = np.concatenate([[0], np.cumsum(dl)]).astype(np.int64)
    full = Bm25Index.from_tokens(torch.from_numpy(tokens).to(dev), torch.from_numpy(doc_ptr).to(dev), vocab)
    dlo, dhi = shard_range(n_docs, world, rank)
    t_lo, t_hi = int(doc_ptr[dlo]), int(doc_ptr[dhi])
    stats = Bm25Stats(n_docs=n_docs, df=full.stats.df, avgdl=full.stats.avgdl)
    shard = Bm25Index.from_tokens(torch.from_numpy(tokens[t_lo:t_hi]).to(dev),
                                  torch.from_numpy(doc_ptr[dlo:dhi + 1] - t_lo).to(dev), vocab, stats=stats, id_base=dlo)
    queries = [rng.integers(0, vocab, 6).tolist() for _ in range(64)]
    ls, li = shard.search(queries, 20)
    gs, gi = exchange_candidates(ls, li)
    ms, mi = ops.merge_candidates(gs, gi, 20)
    ws, wi = full.search(queries, 20)
    torch.cuda.synchronize()
    ok_b = bool(torch.equal(mi, wi)) and bool(torch.equal(ms.view(torch.int32), ws.view(torch.int32)))

    mark(rank, "bm25 shards done")
    # the ShardedBm25Index wrapper and the hybrid (vector + BM25 -> RRF) path == the single-GPU pipeline
    sb = ShardedBm25Index.from_tokens(torch.from_numpy(tokens).to(dev), torch.from_numpy(doc_ptr).to(dev), vocab, n_docs,
                                      stats, rank, world)
    from super_rag_b200.bm25 import prepare_query_terms
    qt = torch.from_numpy(prepare_query_terms(queries, vocab)).to(dev)
    nvec = min(n, n_docs)
    vidx = ShardedVectorIndex(*[t[slice(*shard_range(nvec, world, rank))].contiguous() for t in (cb[:nvec], cs[:nvec])],
                              id_base=shard_range(nvec, world, rank)[0])
    hi_ids, hi_sc = hybrid_search(vidx, sb, Q[:64], qt, 20, 10)
    fvs, fvi = ops.topk_search(cb[:nvec], cs[:nvec], *ops.ingest_rows(Q[:64], "cosine"), 20)
    fbs, fbi = full.search_terms(qt, 20)
    w_ids, w_sc = ops.fuse(torch.stack([fvi, fbi]), None, ops.FUSE_RRF, 10, rank_const=60)
    torch.cuda.synchronize()
    ok_b = ok_b and bool(torch.equal(hi_ids, w_ids)) and bool(torch.equal(hi_sc, w_sc))

    # the same on an index large enough for the two-pass warp-granular path on every shard (>= 32 blocks per shard)
    n_big, vocab_b = 540_000 * world, 400
    rngb = np.random.default_rng(5)
    dlb = rngb.integers(4, 16, n_big)
    pb = 1.0 / np.arange(1, vocab_b + 1)
    tok_b = rngb.choice(vocab_b, size=int(dlb.sum()), p=pb / pb.sum()).astype(np.int64)
    ptr_b = np.concatenate([[0], np.cumsum(dlb)]).astype(np.int64)
    full_b = Bm25Index.from_tokens(torch.from_numpy(tok_b).to(dev), torch.from_numpy(ptr_b).to(dev), vocab_b)
    stats_b = Bm25Stats(n_docs=n_big, df=full_b.stats.df, avgdl=full_b.stats.avgdl)
    sb_b = ShardedBm25Index.from_tokens(torch.from_numpy(tok_b).to(dev), torch.from_numpy(ptr_b).to(dev), vocab_b, n_big,
                                        stats_b, rank, world)
    q_b = [rngb.integers(0, vocab_b, 6).tolist() for _ in range(96)]
    qt_b = torch.from_numpy(prepare_query_terms(q_b, vocab_b)).to(dev)
    bs_, bi_ = sb_b.search_terms(qt_b, 50)
    ws_, wi_ = full_b.search_terms(qt_b, 50)
    torch.cuda.synchronize()
    ok_b = ok_b and bool(torch.equal(bi_, wi_)) and bool(torch.equal(bs_.view(torch.int32), ws_.view(torch.int32)))

    mark(rank, "hybrid and two-pass bm25 shards done")
    flags = torch.tensor([int(ok_v), int(ok_b)], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"world={world} vector_sharded_equals_single={bool(flags[0].item())} "
              f"bm25_and_hybrid_sharded_equals_single={bool(flags[1].item())} cuda_graph_step={graph_used} p2p_exchange={p2p_used}", flush=True)
        if flags[0].item() and flags[1].item():
            print("PARITY OK", flush=True)
    dist.destroy_process_group()
    if not (flags[0].item() and flags[1].item()):
        sys.exit(1)


if __name__ == "__main__":
    main()

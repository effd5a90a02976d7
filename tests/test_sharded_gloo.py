"""N > 1 host logic on CPU: two gloo ranks each own a row shard, produce local top-k candidates with
the oracle (standing in for K1, which needs a GPU), exchange them with the SAME
``sharded.exchange_candidates`` the NCCL path uses, and the merged result must equal the single-shard
answer bit for bit. The K4 merge kernel itself is checked on the GPU (tests/test_gpu_bm25_fuse.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from super_rag_b200.sharded import exchange_candidates, shard_range

N, D, B, K = 3001, 64, 9, 7


def _data():
    rng = np.random.default_rng(11)
    C = oracle.bf16_round(rng.standard_normal((N, D)).astype(np.float32))
    Q = oracle.bf16_round(rng.standard_normal((B, D)).astype(np.float32))
    C[2000] = C[5]          # a cross-shard exact tie: lower id must win
    Q[0] = C[5]
    return C, Q


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        C, Q = _data()
        lo, hi = shard_range(N, world, rank)
        s, i = oracle.exact_topk(C[lo:hi], Q, K, "cosine", id_base=lo)
        gs, gi = exchange_candidates(torch.from_numpy(s), torch.from_numpy(i))
        assert gs.shape == (world, B, K) and gi.shape == (world, B, K)
        ms, mi = oracle.merge_shard_candidates(gs.numpy(), gi.numpy(), K)
        np.save(os.path.join(out_dir, f"s{rank}.npy"), ms)
        np.save(os.path.join(out_dir, f"i{rank}.npy"), mi)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(180)
def test_two_rank_exchange_and_merge_equals_single_shard(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    C, Q = _data()
    want_s, want_i = oracle.exact_topk(C, Q, K, "cosine")
    for r in range(world):
        got_s = np.load(tmp_path / f"s{r}.npy")
        got_i = np.load(tmp_path / f"i{r}.npy")
        assert (got_i == want_i).all(), f"rank {r}: merged ids differ from the single-shard oracle"
        assert (got_s.view(np.uint32) == want_s.view(np.uint32)).all()
    assert want_i[0, 0] == 5 and want_i[0, 1] == 2000

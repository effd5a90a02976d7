"""Row-sharded corpus across the GPUs of one box: one process per GPU, rank r owns a
contiguous row range, every rank scans its shard with K1, the per-shard (score, id) candidates
are all-gathered over NCCL/NVLink and every rank runs the K4 merge kernel.

The only exchange step is that all-gather: 12 * B * k bytes per rank (120 KiB at B=1024, k=10) —
latency-bound, so scores and ids travel in ONE packed int64 buffer and one collective.
The reference has no collective to compare with.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous row range of ``rank``: [r*N/g, (r+1)*N/g) with integer floor."""
    return (rank * n_rows) // world, ((rank + 1) * n_rows) // world


def pack_candidates(scores: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """(fp32 [B,k], int64 [B,k]) -> int64 [2,B,k]: plane 0 holds the score bits, plane 1 the ids."""
    out = torch.empty((2,) + tuple(ids.shape), dtype=torch.int64, device=ids.device)
    out[0] = scores.contiguous().view(torch.int32).to(torch.int64)
    out[1] = ids
    return out


def unpack_candidates(packed: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """int64 [g,2,B,k] -> (scores fp32 [g,B,k], ids int64 [g,B,k])."""
    scores = packed[:, 0].to(torch.int32).contiguous().view(torch.float32)
    ids = packed[:, 1].contiguous()
    return scores, ids


def exchange_candidates(scores: torch.Tensor, ids: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather every rank's [B,k] candidates -> ([g,B,k] scores, [g,B,k] ids) on every rank.
    Works on NCCL (device tensors) and gloo (CPU tensors: the world_size-2 CPU tests)."""
    world = dist.get_world_size(group)
    packed = pack_candidates(scores, ids)
    # concatenation along dim 0 is the one output layout both NCCL and gloo accept
    gathered = torch.empty((world * packed.shape[0],) + tuple(packed.shape[1:]), dtype=torch.int64,
                           device=packed.device)
    dist.all_gather_into_tensor(gathered, packed, group=group)
    return unpack_candidates(gathered.view((world,) + tuple(packed.shape)))


class ShardedVectorIndex:
    """This rank's shard of a row-sharded corpus plus the exchange + merge step."""

    def __init__(self, rows_bf16: torch.Tensor, row_scale: torch.Tensor, id_base: int, metric="cosine", group=None):
        from . import ops
        self.ops = ops
        # the kernels want 16-byte aligned bases; a row slice of a larger tensor need not be
        if rows_bf16.data_ptr() % 16 or not rows_bf16.is_contiguous():
            rows_bf16 = rows_bf16.clone(memory_format=torch.contiguous_format)
        if row_scale.data_ptr() % 16 or not row_scale.is_contiguous():
            row_scale = row_scale.clone(memory_format=torch.contiguous_format)
        self.rows, self.row_scale, self.id_base = rows_bf16, row_scale, int(id_base)
        self.metric = ops.metric_code(metric)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self._ws: Optional[torch.Tensor] = None

    def local_topk(self, q_bf16: torch.Tensor, q_scale: torch.Tensor, k: int):
        need = self.ops.topk_workspace_bytes(self.rows.shape[0], min(q_bf16.shape[0], self.ops.MAX_QUERIES_PER_CALL), k, self.rows.shape[1])
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.rows.device)
        return self.ops.topk_search(self.rows, self.row_scale, q_bf16, q_scale, k, id_base=self.id_base,
                                    workspace=self._ws)

    def search_device(self, q_bf16: torch.Tensor, q_scale: torch.Tensor, k: int):
        """Queries already resident (bf16 + multipliers, replicated on every rank)."""
        s, i = self.local_topk(q_bf16, q_scale, k)
        if self.world == 1:
            return s, i
        gs, gi = exchange_candidates(s, i, self.group)
        return self.ops.merge_candidates(gs, gi, k)

    def search_batch(self, queries: torch.Tensor, k: int):
        """Public API: fp32/bf16 queries on host (pinned) or device -> global top-k on this rank's device."""
        q = queries.to(self.rows.device, non_blocking=True)
        if q.dtype not in (torch.float32, torch.bfloat16):
            q = q.to(torch.float32)
        qb, qs = self.ops.ingest_rows(q.contiguous(), self.metric)
        return self.search_device(qb, qs, k)


class ShardedBm25Index:
    """This rank's document range of a BM25 index built with GLOBAL statistics (df, N, avgdl -> idf), so a
    shard's scores equal the single-index scores bit for bit; search = per-shard K2, one
    all-gather of the (score, id) candidates, K4 merge."""

    def __init__(self, shard_index, group=None):
        from . import ops
        self.ops = ops
        self.index = shard_index                      # bm25.Bm25Index with id_base = first document of the range
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1

    @classmethod
    def from_tokens(cls, tokens: torch.Tensor, doc_ptr: torch.Tensor, vocab: int, n_docs_total: int, stats, rank: int,
                    world: int, group=None) -> "ShardedBm25Index":
        """`tokens` / `doc_ptr` describe the WHOLE collection on every rank (host or device); this rank keeps
        documents shard_range(n_docs_total, world, rank). `stats` = bm25.Bm25Stats of the whole collection."""
        from .bm25 import Bm25Index
        lo, hi = shard_range(n_docs_total, world, rank)
        t_lo, t_hi = int(doc_ptr[lo].item()), int(doc_ptr[hi].item())
        return cls(Bm25Index.from_tokens(tokens[t_lo:t_hi], doc_ptr[lo:hi + 1] - t_lo, vocab, stats=stats, id_base=lo),
                   group)

    def search_terms(self, q_terms: torch.Tensor, k: int):
        s, i = self.index.search_terms(q_terms, k)
        if self.world == 1:
            return s, i
        gs, gi = exchange_candidates(s, i, self.group)
        return self.ops.merge_candidates(gs, gi, k)


def hybrid_search(vec: ShardedVectorIndex, bm25: ShardedBm25Index, queries: torch.Tensor, q_terms: torch.Tensor,
                  k_recall: int, k_out: int, rank_const: int = 60):
    """C4 on shards: GLOBAL vector top-k_recall and GLOBAL BM25 top-k_recall (each: local scan, all-gather,
    merge), then RRF on every rank — the ranks RRF sums are positions in the global lists, so the fusion must
    come after both merges. Returns (ids int64 [B, k_out], scores float64 [B, k_out])."""
    vs, vi = vec.search_batch(queries, k_recall)
    bs, bi = bm25.search_terms(q_terms, k_recall)
    return vec.ops.fuse(torch.stack([vi, bi]), None, vec.ops.FUSE_RRF, k_out, rank_const=rank_const)

"""Row-sharded corpus across the GPUs of one box: one process per GPU, rank r owns a
contiguous row range, every rank scans its shard with K1, the per-shard (score, id) candidates
are all-gathered over NCCL/NVLink and every rank runs the K4 merge kernel.

The only exchange step is that all-gather: 12 * B * k bytes per rank (120 KiB at B=1024, k=10) —
latency-bound, so scores and ids travel in ONE packed int64 buffer and one collective.
The reference has no collective to compare with.
"""
from __future__ import annotations

import threading
from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous row range of ``rank``: [r*N/g, (r+1)*N/g) with integer floor."""
    return (rank * n_rows) // world, ((rank + 1) * n_rows) // world


KEY_ID_LIMIT = 0xFFFFFFFE   # exchange keys carry 32-bit global rows (srag_topk_search_keys)
_M32 = 0xFFFFFFFF


def pack_keys_torch(scores: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """(fp32 [..], int64 [..]) -> int64 [..] exchange keys (uint64 bit patterns): orderable(score) << 32 | (2^32-1 - id),
    0 for an empty slot (id < 0). Plain torch integer ops, any device: the host-side statement of what
    `topk_finalize_kernel` / `srag_pack_keys` emit; the world_size-2 gloo tests run it on CPU tensors."""
    bits = scores.contiguous().view(torch.int32).to(torch.int64) & _M32
    bits = torch.where(bits == 0x80000000, torch.zeros_like(bits), bits)     # -0.0 -> +0.0, as srag_pack_key does
    orderable = torch.where((bits >> 31) != 0, (~bits) & _M32, bits | 0x80000000)
    hi = orderable - ((orderable >> 31) << 32)            # as a signed 32-bit value, so that << 32 cannot overflow
    key = (hi << 32) | ((_M32 - ids) & _M32)
    return torch.where(ids >= 0, key, torch.zeros_like(key))


def unpack_keys_torch(keys: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Inverse of pack_keys_torch: int64 keys -> (fp32 scores, int64 ids); empty -> (-inf, -1)."""
    orderable = (keys >> 32) & _M32
    bits = torch.where((orderable >> 31) != 0, orderable & 0x7FFFFFFF, (~orderable) & _M32)
    bits = bits - ((bits >> 31) << 32)
    scores = bits.to(torch.int32).view(torch.float32)
    ids = _M32 - (keys & _M32)
    empty = keys == 0
    return (torch.where(empty, torch.full_like(scores, float("-inf")), scores),
            torch.where(empty, torch.full_like(ids, -1), ids))


def gather_keys(keys: torch.Tensor, group=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All-gather every rank's [B,k] exchange keys -> [g,B,k] on every rank: ONE collective of 8 bytes per candidate
    (80 KiB per rank at B = 1024, k = 10). NCCL for device tensors, gloo for the CPU tests."""
    world = dist.get_world_size(group)
    B, k = keys.shape
    if out is None:
        out = torch.empty((world, B, k), dtype=torch.int64, device=keys.device)
    dist.all_gather_into_tensor(out.view(world * B, k), keys, group=group)
    return out


def exchange_candidates(scores: torch.Tensor, ids: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather every rank's [B,k] candidates -> ([g,B,k] scores, [g,B,k] ids) on every rank.
    Works on NCCL (device tensors) and gloo (CPU tensors: the world_size-2 CPU tests). Ids beyond the 32-bit key
    range travel as two planes (score bits, ids) instead."""
    world = dist.get_world_size(group)
    if int(ids.max().item() if ids.numel() else 0) < KEY_ID_LIMIT:
        if ids.is_cuda:
            from . import ops
            keys = ops.pack_keys(scores.contiguous(), ids.contiguous())
        else:
            keys = pack_keys_torch(scores, ids)
        return unpack_keys_torch(gather_keys(keys, group))
    packed = torch.empty((2,) + tuple(ids.shape), dtype=torch.int64, device=ids.device)
    packed[0] = scores.contiguous().view(torch.int32).to(torch.int64)
    packed[1] = ids
    gathered = torch.empty((world * 2,) + tuple(ids.shape), dtype=torch.int64, device=ids.device)
    dist.all_gather_into_tensor(gathered, packed, group=group)
    gathered = gathered.view((world, 2) + tuple(ids.shape))
    return gathered[:, 0].to(torch.int32).contiguous().view(torch.float32), gathered[:, 1].contiguous()


class P2pExchange:
    """The symmetric buffer behind `ops.exchange_merge_p2p`: keys uint64 [2][world][cap_keys] + flags uint32 [world][cap_q]
    in ONE torch symmetric-memory allocation per rank, rendezvoused over the process group so that every rank can store
    into every other rank's copy through NVLink. One instance per (index, group); steps are numbered by `epoch`."""

    CAP_Q = 1024           # queries per step (ops.MAX_QUERIES_PER_CALL)
    CAP_K = 128            # SRAG_MAX_TOPK

    def __init__(self, device: torch.device, group=None):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.cap_q, self.cap_keys = self.CAP_Q, self.CAP_Q * self.CAP_K
        key_bytes = 2 * self.world * self.cap_keys * 8
        flag_bytes = self.world * self.cap_q * 4
        self.buf = symm.empty(key_bytes + flag_bytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        pg = group if group is not None else dist.group.WORLD
        self.handle = symm.rendezvous(self.buf, pg)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.key_ptrs = (C.c_uint64 * self.world)(*ptrs)
        self.flag_ptrs = (C.c_uint64 * self.world)(*[p + key_bytes for p in ptrs])
        self.epoch = 0
        dist.barrier(group)   # every rank has zeroed its flags before anyone's first step

    def step(self, ops, local_keys: torch.Tensor, k: int):
        if local_keys.shape[0] > self.cap_q or k > self.CAP_K:
            raise ValueError(f"batch {tuple(local_keys.shape)} exceeds the exchange buffer ({self.cap_q} x {self.CAP_K})")
        self.epoch += 1
        return ops.exchange_merge_p2p(local_keys, self.key_ptrs, self.flag_ptrs, self.world, self.rank, k, self.epoch,
                                      self.cap_keys, self.cap_q)


class ShardedVectorIndex:
    """This rank's shard of a row-sharded corpus plus the exchange + merge step.

    One step on N > 1 GPUs is: workspace memset, K1 (fused GEMM + top-k), the finalize kernel — which writes the
    per-shard candidates as 64-bit exchange keys STRAIGHT INTO the all-gather send buffer —, one NCCL all-gather,
    and the K4 merge kernel reading the gathered keys. Round 1 went through six torch elementwise launches to pack
    and unpack (scores widened to int64, 16 bytes per candidate) around the collective; they cost more than the
    collective itself. With `SRAG_SHARDED_GRAPH=1` the whole step is captured once per (batch, k) into a CUDA graph
    and replayed (queries then live in fixed buffers owned by the index)."""

    def __init__(self, rows_bf16: torch.Tensor, row_scale: torch.Tensor, id_base: int, metric="cosine", group=None):
        import os
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
        self._ws_event = None
        self._lock = threading.Lock()   # the scan workspace belongs to the index: one search at a time
        # exchange keys need every shard's global rows below 2^32 - 1; all ranks must take the same branch
        fits = self.id_base >= 0 and self.id_base + self.rows.shape[0] < KEY_ID_LIMIT
        if self.world > 1:
            flag = torch.tensor([int(fits)], dtype=torch.int32, device=self.rows.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            fits = bool(flag.item())
        self.use_keys = fits
        self.use_graph = self.world > 1 and fits and os.environ.get("SRAG_SHARDED_GRAPH", "0") == "1"
        # SRAG_SHARDED_P2P=1: exchange + merge in one kernel over peer memory instead of the NCCL all-gather (every rank
        # must set it alike); a failed set-up (no symmetric memory / no P2P access) falls back to NCCL on ALL ranks
        self.p2p: Optional[P2pExchange] = None
        if self.world > 1 and fits and os.environ.get("SRAG_SHARDED_P2P", "0") == "1":
            ok = 1
            try:
                self.p2p = P2pExchange(self.rows.device, group)
            except Exception:
                ok = 0
            flag = torch.tensor([ok], dtype=torch.int32, device=self.rows.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            if not bool(flag.item()):
                self.p2p = None
            if self.p2p is not None:
                self.use_graph = False   # the step number is a kernel argument
        self._bufs = {}     # (B, k) -> (send keys [B,k], gathered keys [g,B,k])
        self._graphs = {}   # (B, k) -> (graph, static q_bf16, static q_scale, out scores, out ids)

    def close(self) -> None:
        """Drop the captured CUDA graphs (they hold NCCL work: destroy_process_group() hangs while one is alive)."""
        if self._graphs:
            torch.cuda.synchronize(self.rows.device)
            self._graphs.clear()

    def local_topk(self, q_bf16: torch.Tensor, q_scale: torch.Tensor, k: int, out_keys: Optional[torch.Tensor] = None):
        capturing = torch.cuda.is_current_stream_capturing()
        with self._lock:
            need = self.ops.topk_workspace_bytes(self.rows.shape[0], min(q_bf16.shape[0], self.ops.MAX_QUERIES_PER_CALL), k, self.rows.shape[1])
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=self.rows.device)
            stream = torch.cuda.current_stream(self.rows.device)
            if self._ws_event is not None and not capturing:
                stream.wait_event(self._ws_event)
            out = self.ops.topk_search(self.rows, self.row_scale, q_bf16, q_scale, k, id_base=self.id_base,
                                       workspace=self._ws, out_keys=out_keys, keys_only=out_keys is not None)
            if not capturing:
                self._ws_event = torch.cuda.Event()
                self._ws_event.record(stream)
        return out

    def _step(self, q_bf16: torch.Tensor, q_scale: torch.Tensor, k: int):
        """local scan -> exchange -> merge, on the current stream."""
        if not self.use_keys:
            s, i = self.local_topk(q_bf16, q_scale, k)
            gs, gi = exchange_candidates(s, i, self.group)
            return self.ops.merge_candidates(gs, gi, k)
        B = q_bf16.shape[0]
        bufs = self._bufs.get((B, k))
        if bufs is None:
            dev = self.rows.device
            bufs = (torch.empty(B, k, dtype=torch.int64, device=dev), torch.empty(self.world, B, k, dtype=torch.int64, device=dev))
            self._bufs[(B, k)] = bufs
        send, recv = bufs
        self.local_topk(q_bf16, q_scale, k, out_keys=send)
        if self.p2p is not None:
            return self.p2p.step(self.ops, send, k)
        gather_keys(send, self.group, out=recv)
        return self.ops.merge_keys(recv, k)

    def search_device(self, q_bf16: torch.Tensor, q_scale: torch.Tensor, k: int):
        """Queries already resident (bf16 + multipliers, replicated on every rank)."""
        if self.world == 1:
            return self.local_topk(q_bf16, q_scale, k)
        if not self.use_graph:
            return self._step(q_bf16, q_scale, k)
        return self._replay(q_bf16, q_scale, k)

    def _replay(self, q_bf16: torch.Tensor, q_scale: torch.Tensor, k: int):
        """CUDA-graph path: the step is captured once per (B, k) over fixed query buffers; later calls copy the
        queries in (a no-op when the caller already wrote them there, see `query_buffers`) and replay. The returned
        tensors are the graph's own outputs: consume them before the next search of the same shape."""
        key = (q_bf16.shape[0], k)
        ent = self._graphs.get(key)
        if ent is None:
            sq, ss = q_bf16.clone(), q_scale.clone()
            self._step(sq, ss, k)                       # warm-up outside capture: workspace, buffers, NCCL communicator
            torch.cuda.current_stream(self.rows.device).synchronize()
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    out_s, out_i = self._step(sq, ss, k)
            except Exception:
                self.use_graph = False                  # capture is an optimisation, never a requirement
                return self._step(q_bf16, q_scale, k)
            ent = (g, sq, ss, out_s, out_i)
            self._graphs[key] = ent
        g, sq, ss, out_s, out_i = ent
        if q_bf16.data_ptr() != sq.data_ptr():
            sq.copy_(q_bf16)
        if q_scale.data_ptr() != ss.data_ptr():
            ss.copy_(q_scale)
        g.replay()
        return out_s, out_i

    def submit_batch(self, queries: torch.Tensor, k: int, out_scores: Optional[torch.Tensor] = None,
                     out_ids: Optional[torch.Tensor] = None) -> PendingSearch:
        """`search_batch` without waiting: the host-to-device copy of the queries goes to a copy stream of the index (so
        it overlaps the scan of the batch submitted before), the scan, exchange and merge to the current stream, and the
        results are copied into `out_scores` / `out_ids` (pinned host tensors [B, k] fp32 / int64) when given. A caller
        that keeps two batches in flight — submit i + 1, then wait for i — never leaves the GPU idle between batches;
        every batch still pays its own copies."""
        dev = self.rows.device
        cur = torch.cuda.current_stream(dev)
        if queries.device != dev:
            if getattr(self, "_copy_stream", None) is None:
                self._copy_stream = torch.cuda.Stream(dev)
            with torch.cuda.stream(self._copy_stream):
                q = queries.to(dev, non_blocking=True)
            q.record_stream(cur)
            cur.wait_stream(self._copy_stream)
        else:
            q = queries
        s, i = self.search_batch(q, k)
        done = torch.cuda.Event()
        if out_scores is None:
            done.record(cur)
            return PendingSearch(s, i, done, keep=(q,))
        if self.use_graph and self._graphs:
            # the results are the captured graph's own buffers, rewritten by the next replay: copy them out in stream order
            out_scores.copy_(s, non_blocking=True)
            out_ids.copy_(i, non_blocking=True)
            done.record(cur)
        else:
            # copy-back on its own stream: the next batch's scan does not queue behind two small device-to-host copies
            if getattr(self, "_d2h_stream", None) is None:
                self._d2h_stream = torch.cuda.Stream(dev)
            self._d2h_stream.wait_stream(cur)
            with torch.cuda.stream(self._d2h_stream):
                out_scores.copy_(s, non_blocking=True)
                out_ids.copy_(i, non_blocking=True)
                done.record(self._d2h_stream)
            s.record_stream(self._d2h_stream)
            i.record_stream(self._d2h_stream)
        return PendingSearch(out_scores, out_ids, done, keep=(s, i, q))

    def search_batch(self, queries: torch.Tensor, k: int):
        """Public API: fp32/bf16 queries on host (pinned) or device -> global top-k on this rank's device."""
        q = queries.to(self.rows.device, non_blocking=True)
        if q.dtype not in (torch.float32, torch.bfloat16):
            q = q.to(torch.float32)
        ent = self._graphs.get((q.shape[0], k)) if self.use_graph else None
        if ent is not None and q.dtype == torch.float32:   # cast + normalise straight into the graph's query buffers
            qb, qs = self.ops.ingest_rows(q.contiguous(), self.metric, out_bf16=ent[1], out_scale=ent[2])
        else:
            qb, qs = self.ops.ingest_rows(q.contiguous(), self.metric)
        return self.search_device(qb, qs, k)


class PendingSearch:
    """A search whose kernels and copies have been enqueued (`ShardedVectorIndex.submit_batch`)."""

    def __init__(self, scores, ids, done: "torch.cuda.Event", keep=()):
        self.scores, self.ids, self._done, self._keep = scores, ids, done, keep

    def ready(self) -> bool:
        return self._done.query()

    def wait(self):
        """Block the host until the results are in `scores` / `ids` (the host buffers given to submit_batch, else device
        tensors) and return them."""
        self._done.synchronize()
        self._keep = ()
        return self.scores, self.ids


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
        if self.index.id_base + self.index.n_docs < KEY_ID_LIMIT:      # same on every rank: ranges of one collection
            return self.ops.merge_keys(gather_keys(self.ops.pack_keys(s, i), self.group), k)
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

"""BM25 inverted index resident in HBM + search through K2 (csrc/bm25.cu).

The reference declares a ``fulltext_search`` recall but ships no implementation
(super_rag/nodeflow/runners/merge.py:18-20; super_rag/schema/view_models.py:1043-1047;
docs/design/nodeflow/nodeflow.md:83 names "BM25"). Layout and numeric contract are in
include/srag_b200.h and DESIGN.md §4.2.

Host logic here is limited to: collection statistics (df, avgdl, idf in float64 -> fp32, as the
contract requires) and shaping query term lists. The index build from tokens, impacts, block index, bank order,
scoring and top-k are CUDA kernels.
"""
from __future__ import annotations

import os
import threading
from dataclasses import dataclass
from typing import Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import ops


def idf_from_df(df: np.ndarray, n_docs: int) -> np.ndarray:
    """idf(t) = ln(1 + (N - df + 0.5)/(df + 0.5)), float64 then ONE rounding to fp32."""
    df = np.asarray(df, dtype=np.float64)
    return np.log(1.0 + (float(n_docs) - df + 0.5) / (df + 0.5)).astype(np.float32)


def prepare_query_terms(queries: Sequence[Iterable[int]], vocab: int, max_terms: Optional[int] = None) -> np.ndarray:
    """Unique, in-vocabulary term ids per query in ascending order, -1 padded: int32 [B, T]."""
    cleaned = [sorted({int(t) for t in q if 0 <= int(t) < vocab}) for q in queries]
    width = max([len(c) for c in cleaned] + [1])
    if max_terms is not None:
        if width > max_terms:
            raise ValueError(f"a query has {width} unique terms > max_terms={max_terms}")
        width = max_terms
    out = np.full((len(cleaned), width), -1, dtype=np.int32)
    for b, c in enumerate(cleaned):
        out[b, :len(c)] = c
    return out


@dataclass
class Bm25Stats:
    """Collection statistics; pass the GLOBAL ones to every doc-range shard."""
    n_docs: int
    df: np.ndarray          # int64 [vocab]
    avgdl: float            # float32(sum(dl) / n_docs), division in float64


class Bm25Index:
    """CSR-by-term postings (int32 doc, fp32 impact) + per-term doc-block offsets, all on one GPU."""

    def __init__(self, term_ptr, post_doc, post_imp, block_ptr, n_docs: int, vocab: int, dl, stats: Bm25Stats,
                 id_base: int = 0, bank_ordered: bool = False):
        self.term_ptr, self.post_doc, self.post_imp, self.block_ptr = term_ptr, post_doc, post_imp, block_ptr
        self.n_docs, self.vocab, self.dl, self.stats, self.id_base = n_docs, vocab, dl, stats, id_base
        self.bank_ordered = bank_ordered
        self._ws = None
        self._ws_event = None
        self._lock = threading.Lock()   # the scoring workspace belongs to the index: one search at a time

    def _bank_order(self) -> None:
        """Permute every (term, doc-block) slice for conflict-free accumulation (ops.bm25_bank_order). Scores do not
        depend on the order inside a slice. SRAG_BM25_BANK_ORDER=0 keeps ascending documents (A/B measurements)."""
        if self.bank_ordered or os.environ.get("SRAG_BM25_BANK_ORDER", "1") == "0":
            return
        self.post_doc, self.post_imp = ops.bm25_bank_order(self.block_ptr, self.post_doc, self.post_imp, self.n_docs)
        self.bank_ordered = True

    def canonical_postings(self):
        """(post_doc, post_imp) with ascending documents inside every term — the order the oracle builds."""
        df = self.term_ptr[1:] - self.term_ptr[:-1]
        term = torch.repeat_interleave(torch.arange(self.vocab, device=self.device, dtype=torch.int64), df)
        order = torch.argsort(term * max(1, self.n_docs) + self.post_doc.to(torch.int64))
        return self.post_doc[order], self.post_imp[order]

    @property
    def nnz(self) -> int:
        return int(self.post_doc.numel())

    @property
    def device(self):
        return self.post_doc.device

    # ---- build -----------------------------------------------------------------------------
    @staticmethod
    def postings_from_tokens(tokens: torch.Tensor, doc_ptr: torch.Tensor, vocab: int):
        """(term_ptr int64 [V+1], post_doc int32, post_tf int32, dl int32) from tokens on the device, with the library's
        build kernels (per-document sort + run-length encode, counting, scans, scatter, per-slice sort — ops.
        bm25_build_from_tokens); documents ascend inside a term."""
        term_ptr, post_doc, post_tf, dl, _, _ = ops.bm25_build_from_tokens(tokens, doc_ptr, vocab)
        return term_ptr, post_doc, post_tf, dl

    @classmethod
    def from_postings(cls, term_ptr, post_doc, post_tf, dl, stats: Optional[Bm25Stats] = None, id_base: int = 0,
                      block_ptr=None):
        dev = post_doc.device
        vocab = term_ptr.numel() - 1
        n_docs = dl.numel()
        if stats is None:
            df = (term_ptr[1:] - term_ptr[:-1]).cpu().numpy().astype(np.int64)
            total = float(dl.to(torch.int64).sum().item())
            stats = Bm25Stats(n_docs=n_docs, df=df, avgdl=float(np.float32(total / max(1, n_docs))))
        idf = torch.from_numpy(idf_from_df(stats.df, stats.n_docs)).to(dev)
        # keep the postings 16-byte aligned and padded so 128-bit loads never leave the buffer
        nnz = post_doc.numel()
        pd = torch.zeros(nnz + 4, dtype=torch.int32, device=dev)
        pd[:nnz] = post_doc
        post_doc = pd[:nnz]
        imp = ops.bm25_impacts(term_ptr, post_doc, post_tf, idf, dl, stats.avgdl)
        if block_ptr is None:   # the token build produces the sub-block offsets itself
            block_ptr = ops.bm25_block_index(term_ptr, post_doc, n_docs)
        idx = cls(term_ptr, post_doc, imp, block_ptr, n_docs, vocab, dl, stats, id_base)
        idx._bank_order()
        return idx

    @classmethod
    def from_tokens(cls, tokens: torch.Tensor, doc_ptr: torch.Tensor, vocab: int, stats: Optional[Bm25Stats] = None,
                    id_base: int = 0):
        term_ptr, post_doc, post_tf, dl, _, block_ptr = ops.bm25_build_from_tokens(tokens, doc_ptr, vocab)
        return cls.from_postings(term_ptr, post_doc, post_tf, dl, stats, id_base, block_ptr=block_ptr)

    # ---- persistence (on-disk shard) ------------------------------------
    def save(self, path: str) -> None:
        from . import shardio
        meta = {"n_docs": self.n_docs, "vocab": self.vocab, "id_base": self.id_base, "stats_n_docs": self.stats.n_docs,
                "avgdl": float(self.stats.avgdl), "bank_ordered": bool(self.bank_ordered), "block_docs": ops.bm25_block_docs()}
        shardio.write_shard(path, "bm25", meta, {
            "term_ptr": self.term_ptr, "post_doc": self.post_doc, "post_imp": self.post_imp, "dl": self.dl,
            "df": torch.from_numpy(np.ascontiguousarray(self.stats.df, dtype=np.int64))})

    @classmethod
    def load(cls, path: str, device) -> "Bm25Index":
        """Impacts are stored, so no statistic is recomputed: scores are bit-identical to the saved index.
        Only the per-block offsets are rebuilt (they depend on the library's block size)."""
        from . import shardio
        dev = torch.device(device)
        meta, a, _ = shardio.read_shard(path, device=dev, expect_kind="bm25")
        nnz = a["post_doc"].numel()
        if a["post_imp"].numel() != nnz or a["term_ptr"].numel() != int(meta["vocab"]) + 1 or \
                a["dl"].numel() != int(meta["n_docs"]):
            raise shardio.ShardError(f"{path}: sections do not match the header")
        # 16-byte aligned and padded by one group of 4 postings, as from_postings() lays them out
        pd = torch.zeros(nnz + 4, dtype=torch.int32, device=dev)
        pi = torch.zeros(nnz + 4, dtype=torch.float32, device=dev)
        pd[:nnz] = a["post_doc"]
        pi[:nnz] = a["post_imp"]
        stats = Bm25Stats(n_docs=int(meta["stats_n_docs"]), df=a["df"].cpu().numpy().astype(np.int64),
                          avgdl=float(meta["avgdl"]))
        term_ptr = a["term_ptr"]
        n_docs, vocab = int(meta["n_docs"]), int(meta["vocab"])
        ordered = bool(meta.get("bank_ordered", False))
        idx = cls(term_ptr, pd[:nnz], pi[:nnz], None, n_docs, vocab, a["dl"], stats, int(meta["id_base"]),
                  bank_ordered=ordered)
        if ordered and int(meta.get("block_docs", 0)) != ops.bm25_block_docs():
            # written by a library with another doc-block width: back to ascending documents, then re-ordered below
            d, w = idx.canonical_postings()
            pd[:nnz] = d
            pi[:nnz] = w
            idx.bank_ordered = False
        # The offsets are found by bisection on "doc < j * block_docs", which is monotone along a term as long as its
        # doc-block slices follow one another — true for ascending and for bank-ordered postings of this block width.
        idx.block_ptr = ops.bm25_block_index(term_ptr, idx.post_doc, n_docs)
        idx._bank_order()
        return idx

    # ---- search ----------------------------------------------------------------------------
    def search_terms(self, q_terms: torch.Tensor, k: int, **filter_kw):
        """q_terms int32 [B,T] (unique ascending, -1 padded) on the index device -> (scores, ids).
        filter_kw: doc_tags / q_tag_mask / doc_group / q_group (ops.bm25_search)."""
        if not q_terms.is_cuda:     # host term ids (pinned or not): one small H2D copy
            q_terms = q_terms.to(self.device, non_blocking=True)
        with self._lock:   # ctypes drops the GIL inside srag_*: two searches must not interleave on one workspace
            need = ops.bm25_workspace_bytes(self.n_docs, q_terms.shape[0], k)
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            stream = torch.cuda.current_stream(self.device)
            if self._ws_event is not None:   # ... nor overlap on it from two streams
                stream.wait_event(self._ws_event)
            out = ops.bm25_search(self.block_ptr, self.post_doc, self.post_imp, self.n_docs, q_terms, k,
                                  id_base=self.id_base, workspace=self._ws, **filter_kw)
            self._ws_event = torch.cuda.Event()
            self._ws_event.record(stream)
        return out

    def search(self, queries: Sequence[Iterable[int]], k: int, **filter_kw):
        qt = torch.from_numpy(prepare_query_terms(queries, self.vocab)).to(self.device)
        return self.search_terms(qt, k, **filter_kw)

    def query_posting_bytes(self, q_terms: torch.Tensor) -> int:
        """Algorithmic bytes of a batch: 8 B per posting of every query term."""
        df = (self.term_ptr[1:] - self.term_ptr[:-1])
        t = q_terms.to(torch.int64)
        m = t >= 0
        return int(df[t.clamp(min=0)][m].sum().item()) * 8

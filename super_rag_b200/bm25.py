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
    """CSR-by-term postings as they are SEARCHED, all on one GPU:
      slice_ptr  [V, n_sub+1] (uint32 patterns)  start of every (term, 2048-document sub-block) slice, in groups of 8 postings
      post_off   (uint16 patterns)              byte offset of the document's accumulator inside its 16384-document block's tile
      post_imp   fp32                           impact; 0 marks the padding that rounds every slice up to whole groups
    = 6 bytes per posting. term_ptr (int64 [V+1]) keeps the REAL document frequencies (df = diff). The int32 document ids
    only exist while the index is built; `post_doc` / `canonical_postings` rebuild them on demand (tests, shard conversion)."""

    def __init__(self, term_ptr, slice_ptr, post_off, post_imp, n_docs: int, vocab: int, dl, stats: Bm25Stats,
                 id_base: int = 0, bank_ordered: bool = False):
        self.term_ptr, self.slice_ptr, self.post_off, self.post_imp = term_ptr, slice_ptr, post_off, post_imp
        self.n_docs, self.vocab, self.dl, self.stats, self.id_base = n_docs, vocab, dl, stats, id_base
        self.bank_ordered = bank_ordered
        self._ws = None
        self._ws_event = None
        self._lock = threading.Lock()   # the scoring workspace belongs to the index: one search at a time

    @staticmethod
    def bytes_per_posting() -> int:
        return 6

    @property
    def post_doc(self) -> torch.Tensor:
        """int32 document id of every STORED posting (-1 = padding), in the stored order."""
        return ops.bm25_unpack_docs(self.slice_ptr, self.post_off, self.post_imp, self.n_docs)

    @classmethod
    def _pack(cls, term_ptr, post_doc, post_imp, block_ptr, n_docs, vocab, dl, stats, id_base):
        """Canonical postings -> the searched layout. SRAG_BM25_BANK_ORDER=0 keeps ascending documents inside every slice
        (A/B measurements)."""
        permute = os.environ.get("SRAG_BM25_BANK_ORDER", "1") != "0"
        slice_ptr, post_off, post_imp = ops.bm25_pack(block_ptr, post_doc, post_imp, n_docs, permute=permute)
        return cls(term_ptr, slice_ptr, post_off, post_imp, n_docs, vocab, dl, stats, id_base, bank_ordered=permute)

    def canonical_postings(self):
        """(post_doc int32, post_imp fp32) without padding, ascending documents inside every term — the order the oracle
        builds."""
        doc, term = ops.bm25_unpack_docs(self.slice_ptr, self.post_off, self.post_imp, self.n_docs, with_terms=True)
        keep = doc >= 0
        doc, term, imp = doc[keep], term[keep], self.post_imp[keep]
        order = torch.argsort(term.to(torch.int64) * max(1, self.n_docs) + doc.to(torch.int64))
        return doc[order], imp[order]

    @property
    def nnz(self) -> int:
        """Real postings (padding excluded)."""
        return int(self.term_ptr[-1].item())

    @property
    def nnz_stored(self) -> int:
        return int(self.post_off.numel())

    @property
    def device(self):
        return self.post_off.device

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
        imp = ops.bm25_impacts(term_ptr, post_doc, post_tf, idf, dl, stats.avgdl)
        if block_ptr is None:   # the token build produces the sub-block offsets itself
            block_ptr = ops.bm25_block_index(term_ptr, post_doc, n_docs)
        return cls._pack(term_ptr, post_doc, imp, block_ptr, n_docs, vocab, dl, stats, id_base)

    @classmethod
    def from_tokens(cls, tokens: torch.Tensor, doc_ptr: torch.Tensor, vocab: int, stats: Optional[Bm25Stats] = None,
                    id_base: int = 0):
        term_ptr, post_doc, post_tf, dl, _, block_ptr = ops.bm25_build_from_tokens(tokens, doc_ptr, vocab)
        return cls.from_postings(term_ptr, post_doc, post_tf, dl, stats, id_base, block_ptr=block_ptr)

    # ---- persistence (on-disk shard) ------------------------------------
    def save(self, path: str) -> None:
        """The searched arrays are written as they are (no rebuild on load)."""
        from . import shardio
        meta = {"n_docs": self.n_docs, "vocab": self.vocab, "id_base": self.id_base, "stats_n_docs": self.stats.n_docs,
                "avgdl": float(self.stats.avgdl), "bank_ordered": bool(self.bank_ordered), "layout": "slices-v3",
                "sub_docs": ops.bm25_block_docs(), "abi": 2}
        shardio.write_shard(path, "bm25", meta, {
            "term_ptr": self.term_ptr, "slice_ptr": self.slice_ptr, "post_off": self.post_off, "post_imp": self.post_imp,
            "dl": self.dl, "df": torch.from_numpy(np.ascontiguousarray(self.stats.df, dtype=np.int64))})

    @classmethod
    def load(cls, path: str, device) -> "Bm25Index":
        """Impacts are stored, so no statistic is recomputed: scores are bit-identical to the saved index. A shard in the
        searched layout of this library is used as it is; one written by round 1 (int32 documents) is packed again."""
        from . import shardio
        dev = torch.device(device)
        meta, a, _ = shardio.read_shard(path, device=dev, expect_kind="bm25")
        n_docs, vocab = int(meta["n_docs"]), int(meta["vocab"])
        if a["term_ptr"].numel() != vocab + 1 or a["dl"].numel() != n_docs:
            raise shardio.ShardError(f"{path}: sections do not match the header")
        stats = Bm25Stats(n_docs=int(meta["stats_n_docs"]), df=a["df"].cpu().numpy().astype(np.int64),
                          avgdl=float(meta["avgdl"]))
        term_ptr = a["term_ptr"]
        if meta.get("layout") == "slices-v3" and int(meta.get("sub_docs", 0)) == ops.bm25_block_docs():
            n = a["post_off"].numel()
            n_sub = max(1, -(-n_docs // ops.bm25_block_docs()))
            if a["post_imp"].numel() != n or tuple(a["slice_ptr"].shape) != (vocab, n_sub + 1) or n % 8:
                raise shardio.ShardError(f"{path}: sections do not match the header")
            po = torch.zeros(n + ops.BM25_PAD, dtype=torch.int16, device=dev)
            pi = torch.zeros(n + ops.BM25_PAD, dtype=torch.float32, device=dev)
            po[:n] = a["post_off"]
            pi[:n] = a["post_imp"]
            return cls(term_ptr, a["slice_ptr"].contiguous(), po[:n], pi[:n], n_docs, vocab, a["dl"], stats,
                       int(meta["id_base"]), bank_ordered=bool(meta.get("bank_ordered", False)))
        if "post_doc" not in a:
            raise shardio.ShardError(f"{path}: unknown BM25 shard layout {meta.get('layout')!r}")
        # round-1 shard: int32 documents, ascending or bank-ordered per 2048-document slice. The slice offsets are found
        # by bisection on "doc < j * sub_docs", monotone along a term either way; slices are then sorted back to ascending
        # documents (they are sets) and packed.
        pd, pi = a["post_doc"].contiguous(), a["post_imp"].contiguous()
        if pi.numel() != pd.numel():
            raise shardio.ShardError(f"{path}: sections do not match the header")
        df = term_ptr[1:] - term_ptr[:-1]
        term = torch.repeat_interleave(torch.arange(vocab, device=dev, dtype=torch.int64), df)
        order = torch.argsort(term * max(1, n_docs) + pd.to(torch.int64))
        pd, pi = pd[order].contiguous(), pi[order].contiguous()
        block_ptr = ops.bm25_block_index(term_ptr, pd, n_docs)
        return cls._pack(term_ptr, pd, pi, block_ptr, n_docs, vocab, a["dl"], stats, int(meta["id_base"]))

    # ---- search ----------------------------------------------------------------------------
    def search_terms(self, q_terms: torch.Tensor, k: int, **filter_kw):
        """q_terms int32 [B,T] (unique ascending, -1 padded) on the index device -> (scores, ids).
        filter_kw: doc_tags / q_tag_mask / doc_group / q_group (ops.bm25_search)."""
        if self.nnz_stored == 0:    # no postings (every text tokenised to nothing): nothing can match
            B = q_terms.shape[0]
            return (torch.full((B, k), float("-inf"), dtype=torch.float32, device=self.device),
                    torch.full((B, k), -1, dtype=torch.int64, device=self.device))
        if not q_terms.is_cuda:     # host term ids (pinned or not): one small H2D copy
            q_terms = q_terms.to(self.device, non_blocking=True)
        with self._lock:   # ctypes drops the GIL inside srag_*: two searches must not interleave on one workspace
            need = ops.bm25_workspace_bytes(self.n_docs, q_terms.shape[0], k)
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            stream = torch.cuda.current_stream(self.device)
            if self._ws_event is not None:   # ... nor overlap on it from two streams
                stream.wait_event(self._ws_event)
            out = ops.bm25_search(self.slice_ptr, self.post_off, self.post_imp, self.n_docs, q_terms, k,
                                  id_base=self.id_base, workspace=self._ws, **filter_kw)
            self._ws_event = torch.cuda.Event()
            self._ws_event.record(stream)
        return out

    def search(self, queries: Sequence[Iterable[int]], k: int, **filter_kw):
        qt = torch.from_numpy(prepare_query_terms(queries, self.vocab)).to(self.device)
        return self.search_terms(qt, k, **filter_kw)

    def query_posting_bytes(self, q_terms: torch.Tensor) -> int:
        """Algorithmic bytes of a batch: 8 B per posting of every query term; the index
        itself stores 6 B per posting, so the HBM traffic of the scan is below this figure."""
        df = (self.term_ptr[1:] - self.term_ptr[:-1])
        t = q_terms.to(torch.int64)
        m = t >= 0
        return int(df[t.clamp(min=0)][m].sum().item()) * 8

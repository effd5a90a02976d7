"""Thin torch-tensor front end of the C ABI (include/srag_b200.h). PyTorch supplies device
memory and streams only; every computation below is a libsrag_b200 kernel. No CPU fallback:
tensors must be CUDA tensors and failures raise ``SragError``."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import SragError

METRIC_COSINE, METRIC_IP = 0, 1
FUSE_RRF, FUSE_WEIGHTED = 0, 1
MAX_TOPK, MAX_BM25_K, MAX_LISTS = 128, 256, 8
# batches are scanned 1024 queries per launch: 4 query tiles x 18 slabs keeps threshold sharing (j <= 8) on for k <= 128
MAX_QUERIES_PER_CALL = 1024


def metric_code(metric) -> int:
    if metric in (METRIC_COSINE, "cosine"):
        return METRIC_COSINE
    if metric in (METRIC_IP, "ip", "inner_product", "dot"):
        return METRIC_IP
    raise ValueError(f"unsupported distance {metric!r} (cosine | ip)")


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise SragError("libsrag_b200 operates on CUDA tensors only (no CPU fallback)")


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def device_info() -> Tuple[int, int, int]:
    lib = _lib.load()
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    _lib.check(lib.srag_device_info(C.byref(a), C.byref(b), C.byref(c)), "srag_device_info")
    return a.value, b.value, c.value


# ---- K0 ---------------------------------------------------------------------------------
def ingest_rows(rows: torch.Tensor, metric=METRIC_COSINE, out_bf16: Optional[torch.Tensor] = None,
                out_scale: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """fp32 (cast + multiplier) or bf16 (multiplier only) rows -> (bf16 rows, fp32 row multipliers)."""
    _need_cuda(rows)
    lib = _lib.load()
    rows = rows.contiguous()
    n, dim = rows.shape
    m = metric_code(metric)
    with torch.cuda.device(rows.device):
        scale = out_scale if out_scale is not None else torch.empty(n, dtype=torch.float32, device=rows.device)
        if rows.dtype == torch.bfloat16:
            _lib.check(lib.srag_rows_scale_bf16(_ptr(rows), n, dim, _ptr(scale), m, _stream(rows.device)),
                       "srag_rows_scale_bf16")
            if out_bf16 is not None:
                out_bf16.copy_(rows)
                return out_bf16, scale
            return rows, scale
        if rows.dtype != torch.float32:
            raise SragError(f"rows must be float32 or bfloat16, got {rows.dtype}")
        dst = out_bf16 if out_bf16 is not None else torch.empty(n, dim, dtype=torch.bfloat16, device=rows.device)
        _lib.check(lib.srag_rows_ingest_f32(_ptr(rows), n, dim, _ptr(dst), _ptr(scale), m, _stream(rows.device)),
                   "srag_rows_ingest_f32")
    return dst, scale


# ---- K1 ---------------------------------------------------------------------------------
def topk_workspace_bytes(n_rows: int, n_queries: int, k: int, dim: int = 0) -> int:
    """With `dim` the size includes room for the zero-padded query copy that small / ragged batches use."""
    lib = _lib.load()
    out = C.c_size_t()
    if dim:
        _lib.check(lib.srag_topk_workspace_bytes_ex(n_rows, dim, n_queries, k, C.byref(out)), "srag_topk_workspace_bytes_ex")
    else:
        _lib.check(lib.srag_topk_workspace_bytes(n_rows, n_queries, k, C.byref(out)), "srag_topk_workspace_bytes")
    return out.value


def topk_search(corpus: torch.Tensor, row_scale: torch.Tensor, queries: torch.Tensor, q_scale: torch.Tensor,
                k: int, id_base: int = 0, workspace: Optional[torch.Tensor] = None,
                out_scores: Optional[torch.Tensor] = None, out_ids: Optional[torch.Tensor] = None,
                row_tags: Optional[torch.Tensor] = None, q_tag_mask: Optional[torch.Tensor] = None,
                row_group: Optional[torch.Tensor] = None, q_group: Optional[torch.Tensor] = None,
                out_keys: Optional[torch.Tensor] = None, keys_only: bool = False):
    """Exact top-k of bf16 queries [B,D] against a bf16 corpus [N,D]; returns (scores fp32 [B,k], ids int64 [B,k]).

    `out_keys` (int64 [B,k] holding uint64 bit patterns): also emit the cross-shard exchange keys (srag_topk_search_keys);
    with `keys_only` scores / ids are not written and (None, None) is returned.

    Optional metadata pre-filter (int32 tensors holding 32-bit patterns): row r is eligible for query b iff
    ``row_tags[r] & q_tag_mask[b] != 0`` (when given) and ``q_group[b] < 0 or row_group[r] == q_group[b]`` (when given)."""
    _need_cuda(corpus, row_scale, queries, q_scale)
    if (row_tags is None) != (q_tag_mask is None) or (row_group is None) != (q_group is None):
        raise SragError("row_tags / q_tag_mask and row_group / q_group come in pairs")
    for t, n_expect, what in ((row_tags, corpus.shape[0], "row_tags"), (row_group, corpus.shape[0], "row_group"),
                              (q_tag_mask, queries.shape[0], "q_tag_mask"), (q_group, queries.shape[0], "q_group")):
        if t is not None:
            _need_cuda(t)
            if t.dtype != torch.int32 or t.numel() != n_expect or not t.is_contiguous():
                raise SragError(f"{what} must be a contiguous int32 tensor of {n_expect} elements")
    filtered = row_tags is not None or row_group is not None
    if out_keys is not None:
        _need_cuda(out_keys)
        if filtered:
            raise SragError("exchange keys are not available with the metadata filter")
        if out_keys.dtype != torch.int64 or tuple(out_keys.shape) != (queries.shape[0], k) or not out_keys.is_contiguous():
            raise SragError("out_keys must be a contiguous int64 [B, k] tensor")
    if corpus.dtype != torch.bfloat16 or queries.dtype != torch.bfloat16:
        raise SragError("corpus and queries must be bfloat16 (use ingest_rows)")
    lib = _lib.load()
    B, D = queries.shape
    N = corpus.shape[0]
    if N and corpus.shape[1] != D:
        raise SragError(f"dimension mismatch: corpus {tuple(corpus.shape)} vs queries {tuple(queries.shape)}")
    dev = queries.device
    with torch.cuda.device(dev):
        want_pairs = not (keys_only and out_keys is not None)
        scores = ids = None
        if want_pairs:
            scores = out_scores if out_scores is not None else torch.empty(B, k, dtype=torch.float32, device=dev)
            ids = out_ids if out_ids is not None else torch.empty(B, k, dtype=torch.int64, device=dev)
        for b0 in range(0, B, MAX_QUERIES_PER_CALL):
            nb = min(MAX_QUERIES_PER_CALL, B - b0)
            need = topk_workspace_bytes(N, nb, k, D)
            if workspace is None or workspace.numel() < need:
                workspace = torch.empty(need, dtype=torch.uint8, device=dev)
            q = queries[b0:b0 + nb]
            if filtered:
                _lib.check(lib.srag_topk_search_filtered(
                    _ptr(corpus), _ptr(row_scale), N, D, _ptr(q), _ptr(q_scale[b0:b0 + nb]), nb, k, id_base,
                    _ptr(row_tags), _ptr(row_group), _ptr(q_tag_mask[b0:b0 + nb] if q_tag_mask is not None else None),
                    _ptr(q_group[b0:b0 + nb] if q_group is not None else None), _ptr(scores[b0:b0 + nb]),
                    _ptr(ids[b0:b0 + nb]), _ptr(workspace), workspace.numel(), _stream(dev)), "srag_topk_search_filtered")
                continue
            if out_keys is not None:
                _lib.check(lib.srag_topk_search_keys(
                    _ptr(corpus), _ptr(row_scale), N, D, _ptr(q), _ptr(q_scale[b0:b0 + nb]), nb, k, id_base,
                    _ptr(out_keys[b0:b0 + nb]), _ptr(scores[b0:b0 + nb] if want_pairs else None),
                    _ptr(ids[b0:b0 + nb] if want_pairs else None), _ptr(workspace), workspace.numel(), _stream(dev)),
                    "srag_topk_search_keys")
                continue
            _lib.check(lib.srag_topk_search(_ptr(corpus), _ptr(row_scale), N, D, _ptr(q), _ptr(q_scale[b0:b0 + nb]),
                                            nb, k, id_base, _ptr(scores[b0:b0 + nb]), _ptr(ids[b0:b0 + nb]),
                                            _ptr(workspace), workspace.numel(), _stream(dev)), "srag_topk_search")
    return scores, ids


TOPK_STATIC_SCRATCH_BYTES = 8 << 20


def topk_search_static(corpus: torch.Tensor, row_scale: torch.Tensor, queries: torch.Tensor, q_scale: torch.Tensor, k: int,
                       out_scores: torch.Tensor, out_ids: torch.Tensor, id_base: int = 0) -> None:
    """srag_topk_search_static: the exact top-k scan with the library's own static device scratch (8 MiB) — no workspace,
    no device allocation anywhere in the call (this wrapper allocates nothing either: the caller passes the outputs).
    No counterpart in the reference. Calls on one device share the scratch:
    do not overlap them on the GPU. Raises SragError when the search needs more scratch than the array holds."""
    _need_cuda(corpus, row_scale, queries, q_scale, out_scores, out_ids)
    if corpus.dtype != torch.bfloat16 or queries.dtype != torch.bfloat16:
        raise SragError("corpus and queries must be bfloat16 (use ingest_rows)")
    B, D = queries.shape
    if out_scores.dtype != torch.float32 or out_ids.dtype != torch.int64 or tuple(out_scores.shape) != (B, k) or \
            tuple(out_ids.shape) != (B, k) or not out_scores.is_contiguous() or not out_ids.is_contiguous():
        raise SragError("out_scores fp32 [B, k] and out_ids int64 [B, k], contiguous")
    N = corpus.shape[0]
    if N and corpus.shape[1] != D:
        raise SragError(f"dimension mismatch: corpus {tuple(corpus.shape)} vs queries {tuple(queries.shape)}")
    dev = queries.device
    with torch.cuda.device(dev):
        _lib.check(_lib.load().srag_topk_search_static(_ptr(corpus), _ptr(row_scale), N, D, _ptr(queries), _ptr(q_scale), B, k,
                                                       id_base, _ptr(out_scores), _ptr(out_ids), _stream(dev)),
                   "srag_topk_search_static")


def prof_enable(on: bool):
    _lib.check(_lib.load().srag_prof_enable(1 if on else 0), "srag_prof_enable")


def prof_read():
    """(sum of main-kernel ms, sum of finalize ms, calls) since prof_enable(True)."""
    a, b, n = C.c_double(), C.c_double(), C.c_int()
    _lib.check(_lib.load().srag_prof_read(C.byref(a), C.byref(b), C.byref(n)), "srag_prof_read")
    return a.value, b.value, n.value


# ---- K2 ---------------------------------------------------------------------------------
def bm25_block_docs() -> int:
    return _lib.load().srag_bm25_block_docs()


def bm25_workspace_bytes(n_docs: int, n_queries: int, k: int) -> int:
    out = C.c_size_t()
    _lib.check(_lib.load().srag_bm25_workspace_bytes(n_docs, n_queries, k, C.byref(out)), "srag_bm25_workspace_bytes")
    return out.value


def bm25_impacts(term_ptr, post_doc, post_tf, idf, dl, avgdl: float) -> torch.Tensor:
    _need_cuda(term_ptr, post_doc, post_tf, idf, dl)
    lib = _lib.load()
    nnz = post_doc.numel()
    dev = post_doc.device
    with torch.cuda.device(dev):
        imp = torch.empty(nnz + 4, dtype=torch.float32, device=dev)[:nnz]
        _lib.check(lib.srag_bm25_impacts(_ptr(term_ptr), term_ptr.numel() - 1, _ptr(post_doc), _ptr(post_tf), nnz,
                                         _ptr(idf), _ptr(dl), float(avgdl), _ptr(imp), _stream(dev)),
                   "srag_bm25_impacts")
    return imp


def bm25_block_index(term_ptr, post_doc, n_docs: int) -> torch.Tensor:
    _need_cuda(term_ptr, post_doc)
    lib = _lib.load()
    vocab = term_ptr.numel() - 1
    bd = bm25_block_docs()
    nb = max(1, (n_docs + bd - 1) // bd)
    dev = term_ptr.device
    with torch.cuda.device(dev):
        bp = torch.empty(vocab, nb + 1, dtype=torch.int64, device=dev)
        _lib.check(lib.srag_bm25_block_index(_ptr(term_ptr), _ptr(post_doc), vocab, n_docs, bd, _ptr(bp), _stream(dev)),
                   "srag_bm25_block_index")
    return bp


def bm25_build_from_tokens(tokens: torch.Tensor, doc_ptr: torch.Tensor, vocab: int):
    """CSR postings from tokens with the library's own kernels (srag_bm25_build_plan / _fill; no global sort).
    tokens: integer term ids [n_tokens] on the GPU, doc_ptr int64 [n_docs+1]. Returns (term_ptr int64 [V+1], post_doc int32
    [nnz], post_tf int32 [nnz], dl int32 [n_docs], df int64 [V], block_ptr int64 [V, n_sub+1]); documents ascend inside a term."""
    _need_cuda(tokens, doc_ptr)
    lib = _lib.load()
    dev = tokens.device
    n_docs = doc_ptr.numel() - 1
    bd = bm25_block_docs()
    n_sub = max(1, (n_docs + bd - 1) // bd)
    with torch.cuda.device(dev):
        toks = tokens.to(torch.int32, copy=True).contiguous()   # consumed: becomes the per-document distinct terms
        if toks.numel() == 0:                                    # documents without a single token: keep the pointers non-null
            toks = torch.zeros(1, dtype=torch.int32, device=dev)
        dp = doc_ptr.to(torch.int64).contiguous()
        uniq_tf = torch.empty_like(toks)
        doc_uniq = torch.empty(max(1, n_docs), dtype=torch.int32, device=dev)
        dl = torch.empty(max(1, n_docs), dtype=torch.int32, device=dev)
        df = torch.empty(vocab, dtype=torch.int64, device=dev)
        term_ptr = torch.empty(vocab + 1, dtype=torch.int64, device=dev)
        block_ptr = torch.empty(vocab, n_sub + 1, dtype=torch.int64, device=dev)
        err = torch.zeros(1, dtype=torch.int32, device=dev)
        st = _stream(dev)
        _lib.check(lib.srag_bm25_build_plan(_ptr(toks), _ptr(dp), n_docs, vocab, _ptr(uniq_tf), _ptr(doc_uniq), _ptr(dl),
                                            _ptr(df), _ptr(term_ptr), _ptr(block_ptr), _ptr(err), st), "srag_bm25_build_plan")
        nnz, flag = int(term_ptr[vocab].item()), int(err.item())   # the one host sync of the build
        if flag == 1:
            raise SragError("srag_bm25_build_plan: a document has more than 57344 tokens")
        if flag == 2:
            raise SragError(f"srag_bm25_build_plan: a term id is outside [0, {vocab})")
        post_doc = torch.zeros(nnz + 4, dtype=torch.int32, device=dev)[:nnz]   # padded: 128-bit loads stay inside
        post_tf = torch.zeros(nnz + 4, dtype=torch.int32, device=dev)[:nnz]
        fill = torch.empty(max(1, vocab * n_sub), dtype=torch.int32, device=dev)
        if nnz:   # nothing to fill for a collection whose texts all tokenise to nothing
            _lib.check(lib.srag_bm25_build_fill(_ptr(toks), _ptr(uniq_tf), _ptr(dp), _ptr(doc_uniq), n_docs, vocab,
                                                _ptr(block_ptr), _ptr(fill), _ptr(post_doc), _ptr(post_tf), st),
                       "srag_bm25_build_fill")
    return term_ptr, post_doc, post_tf, dl[:n_docs], df, block_ptr


BM25_PAD = 8   # entries of padding behind the searched posting arrays (bulk copies and 128-bit loads stay inside)


def bm25_pack(block_ptr, post_doc, post_imp, n_docs: int, permute: bool = True):
    """Canonical postings -> the searched layout (srag_bm25_pack_plan / _fill): every (term, sub-block) slice starts on a
    group of 8 postings and is padded to whole groups with zero-impact postings, documents become 16-bit tile offsets and
    (permute) the postings of every slice are laid out for conflict-free shared-memory accumulation.
    Returns (slice_ptr int32-typed [V, n_sub+1] holding uint32 group indexes, post_off int16-typed (uint16 patterns),
    post_imp fp32), the two posting arrays padded behind their end; scores are unchanged. One host sync (the padded size)."""
    _need_cuda(block_ptr, post_doc, post_imp)
    lib = _lib.load()
    dev = post_doc.device
    vocab = block_ptr.shape[0]
    with torch.cuda.device(dev):
        scratch = torch.empty(2 * vocab + 1, dtype=torch.int64, device=dev)
        slice_ptr = torch.empty(block_ptr.shape, dtype=torch.int32, device=dev)
        st = _stream(dev)
        _lib.check(lib.srag_bm25_pack_plan(_ptr(block_ptr), vocab, n_docs, _ptr(scratch), _ptr(slice_ptr), st), "srag_bm25_pack_plan")
        groups = int(scratch[2 * vocab].item())
        if groups >= 1 << 32:
            raise SragError(f"{groups} groups of 8 postings do not fit the 32-bit slice table: shard the index")
        n_pad = 8 * groups
        po = torch.zeros(n_pad + BM25_PAD, dtype=torch.int16, device=dev)[:n_pad]
        pi = torch.zeros(n_pad + BM25_PAD, dtype=torch.float32, device=dev)[:n_pad]
        if post_doc.numel():
            _lib.check(lib.srag_bm25_pack_fill(_ptr(block_ptr), _ptr(slice_ptr), vocab, n_docs, _ptr(post_doc), _ptr(post_imp),
                                               _ptr(po), _ptr(pi), 1 if permute else 0, st), "srag_bm25_pack_fill")
    return slice_ptr, po, pi


def bm25_unpack_docs(slice_ptr, post_off, post_imp, n_docs: int, with_terms: bool = False):
    """int32 document id of every stored posting (-1 = padding), in the stored order (srag_bm25_unpack_docs); with
    `with_terms` also the term of every posting. Not on the search path."""
    _need_cuda(slice_ptr, post_off, post_imp)
    n = post_off.numel()
    dev = post_off.device
    with torch.cuda.device(dev):
        doc = torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev)[:n]
        term = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)[:n] if with_terms else None
        if n:
            _lib.check(_lib.load().srag_bm25_unpack_docs(_ptr(slice_ptr), slice_ptr.shape[0], n_docs, _ptr(post_off), _ptr(post_imp),
                                                         _ptr(doc), _ptr(term), _stream(dev)), "srag_bm25_unpack_docs")
    return (doc, term) if with_terms else doc


def bm25_search(slice_ptr, post_off, post_imp, n_docs: int, q_terms: torch.Tensor, k: int, id_base: int = 0,
                workspace: Optional[torch.Tensor] = None, doc_tags: Optional[torch.Tensor] = None,
                q_tag_mask: Optional[torch.Tensor] = None, doc_group: Optional[torch.Tensor] = None,
                q_group: Optional[torch.Tensor] = None):
    """q_terms int32 [B, T]: unique ascending term ids, -1 padded. Optional metadata pre-filter as in
    topk_search (int32 tensors: doc_tags / doc_group [n_docs], q_tag_mask / q_group [B])."""
    _need_cuda(slice_ptr, post_off, post_imp, q_terms)
    if slice_ptr.dtype != torch.int32 or post_off.dtype != torch.int16 or post_imp.dtype != torch.float32 or \
            post_off.numel() != post_imp.numel():
        raise SragError("bm25_search wants the searched layout of ops.bm25_pack: slice_ptr (32-bit), post_off (16-bit), post_imp fp32")
    if (doc_tags is None) != (q_tag_mask is None) or (doc_group is None) != (q_group is None):
        raise SragError("doc_tags / q_tag_mask and doc_group / q_group come in pairs")
    for t, n_expect, what in ((doc_tags, n_docs, "doc_tags"), (doc_group, n_docs, "doc_group"),
                              (q_tag_mask, q_terms.shape[0], "q_tag_mask"), (q_group, q_terms.shape[0], "q_group")):
        if t is not None:
            _need_cuda(t)
            if t.dtype != torch.int32 or t.numel() != n_expect or not t.is_contiguous():
                raise SragError(f"{what} must be a contiguous int32 tensor of {n_expect} elements")
    lib = _lib.load()
    B, T = q_terms.shape
    dev = q_terms.device
    vocab = slice_ptr.shape[0]
    with torch.cuda.device(dev):
        out = C.c_size_t()
        _lib.check(lib.srag_bm25_workspace_bytes(n_docs, B, k, C.byref(out)), "srag_bm25_workspace_bytes")
        if workspace is None or workspace.numel() < out.value:
            workspace = torch.empty(out.value, dtype=torch.uint8, device=dev)
        scores = torch.empty(B, k, dtype=torch.float32, device=dev)
        ids = torch.empty(B, k, dtype=torch.int64, device=dev)
        if doc_tags is not None or doc_group is not None:
            _lib.check(lib.srag_bm25_search_filtered(
                _ptr(slice_ptr), _ptr(post_off), _ptr(post_imp), vocab, n_docs, _ptr(q_terms.contiguous()), B, T, k,
                id_base, _ptr(doc_tags), _ptr(doc_group), _ptr(q_tag_mask), _ptr(q_group), _ptr(scores), _ptr(ids),
                _ptr(workspace), workspace.numel(), _stream(dev)), "srag_bm25_search_filtered")
        else:
            _lib.check(lib.srag_bm25_search(_ptr(slice_ptr), _ptr(post_off), _ptr(post_imp), vocab, n_docs,
                                            _ptr(q_terms.contiguous()), B, T, k, id_base, _ptr(scores), _ptr(ids),
                                            _ptr(workspace), workspace.numel(), _stream(dev)), "srag_bm25_search")
    return scores, ids


# ---- K5 -----------------------------------------------------------------------------------
MAX_MMR = 256


def mmr_rerank(corpus: torch.Tensor, row_scale: torch.Tensor, queries: torch.Tensor, q_scale: torch.Tensor,
               cand_ids: torch.Tensor, mmr_lambda: float = 0.5, min_score: float = -2.0, id_base: int = 0):
    """Reference MMR (search_utils.py:1884-1922) over candidate rows of the device corpus.
    cand_ids int64 [B, m] (-1 = none) -> (ids int64 [B, m], scores fp32 [B, m]) by score desc, -1 / -inf padded."""
    _need_cuda(corpus, row_scale, queries, q_scale, cand_ids)
    if corpus.dtype != torch.bfloat16 or queries.dtype != torch.bfloat16 or cand_ids.dtype != torch.int64:
        raise SragError("corpus / queries must be bfloat16 and cand_ids int64")
    B, m = cand_ids.shape
    if queries.shape[0] != B or q_scale.numel() != B:
        raise SragError("one query (and q_scale) per candidate list")
    dev = queries.device
    with torch.cuda.device(dev):
        out_ids = torch.empty(B, m, dtype=torch.int64, device=dev)
        out_scores = torch.empty(B, m, dtype=torch.float32, device=dev)
        _lib.check(_lib.load().srag_mmr_rerank(_ptr(corpus), _ptr(row_scale), corpus.shape[0], queries.shape[1],
                                               _ptr(queries.contiguous()), _ptr(q_scale), B, _ptr(cand_ids.contiguous()), m,
                                               id_base, float(mmr_lambda), float(min_score), _ptr(out_ids),
                                               _ptr(out_scores), _stream(dev)), "srag_mmr_rerank")
    return out_ids, out_scores


# ---- K3 / K4 ----------------------------------------------------------------------------
def fuse(ids: torch.Tensor, scores: Optional[torch.Tensor], mode: int, k_out: int, rank_const: int = 1,
         min_score: float = 0.0, weights: Optional[Sequence[float]] = None):
    """ids int64 [L, B, K] (id < 0 = padding), scores fp32 [L, B, K] -> (ids int64 [B,k_out], scores f64 [B,k_out])."""
    _need_cuda(ids, scores)
    lib = _lib.load()
    L, B, K = ids.shape
    dev = ids.device
    ids = ids.contiguous()
    if scores is not None:
        scores = scores.contiguous().to(torch.float32)
    w = None
    if weights is not None:
        w = (C.c_double * L)(*[float(x) for x in weights])
    with torch.cuda.device(dev):
        oi = torch.empty(B, k_out, dtype=torch.int64, device=dev)
        os_ = torch.empty(B, k_out, dtype=torch.float64, device=dev)
        _lib.check(lib.srag_fuse(_ptr(ids), _ptr(scores), L, B, K, mode, rank_const, float(min_score), w, k_out,
                                 _ptr(oi), _ptr(os_), _stream(dev)), "srag_fuse")
    return oi, os_


def merge_candidates(scores: torch.Tensor, ids: torch.Tensor, k: int):
    """scores fp32 [G, B, KK], ids int64 [G, B, KK] -> global top-k (scores [B,k], ids [B,k])."""
    _need_cuda(scores, ids)
    lib = _lib.load()
    G, B, KK = ids.shape
    dev = ids.device
    with torch.cuda.device(dev):
        os_ = torch.empty(B, k, dtype=torch.float32, device=dev)
        oi = torch.empty(B, k, dtype=torch.int64, device=dev)
        _lib.check(lib.srag_merge_candidates(_ptr(scores.contiguous()), _ptr(ids.contiguous()), G, B, KK, k,
                                             _ptr(os_), _ptr(oi), _stream(dev)), "srag_merge_candidates")
    return os_, oi


def merge_keys(keys: torch.Tensor, k: int, out_scores: Optional[torch.Tensor] = None,
               out_ids: Optional[torch.Tensor] = None):
    """keys int64 [G, B, KK] (uint64 exchange-key bit patterns, as all-gathered) -> global top-k (scores [B,k], ids [B,k])."""
    _need_cuda(keys)
    if keys.dtype != torch.int64 or keys.ndim != 3 or not keys.is_contiguous():
        raise SragError("keys must be a contiguous int64 [G, B, KK] tensor")
    G, B, KK = keys.shape
    dev = keys.device
    with torch.cuda.device(dev):
        os_ = out_scores if out_scores is not None else torch.empty(B, k, dtype=torch.float32, device=dev)
        oi = out_ids if out_ids is not None else torch.empty(B, k, dtype=torch.int64, device=dev)
        _lib.check(_lib.load().srag_merge_keys(_ptr(keys), G, B, KK, k, _ptr(os_), _ptr(oi), _stream(dev)), "srag_merge_keys")
    return os_, oi


def pack_keys(scores: torch.Tensor, ids: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(fp32 scores, int64 ids < 2^32 - 1; id < 0 = empty) -> int64 tensor of exchange-key bit patterns, same shape."""
    _need_cuda(scores, ids)
    if scores.dtype != torch.float32 or ids.dtype != torch.int64 or scores.shape != ids.shape:
        raise SragError("pack_keys wants fp32 scores and int64 ids of one shape")
    dev = ids.device
    with torch.cuda.device(dev):
        keys = out if out is not None else torch.empty(ids.shape, dtype=torch.int64, device=dev)
        _lib.check(_lib.load().srag_pack_keys(_ptr(scores.contiguous()), _ptr(ids.contiguous()), ids.numel(), _ptr(keys),
                                              _ptr(None), _stream(dev)), "srag_pack_keys")
    return keys


def exchange_merge_p2p(local_keys: torch.Tensor, peer_key_ptrs, peer_flag_ptrs, world: int, rank: int, k: int, epoch: int,
                       cap_keys: int, cap_q: int, out_scores: Optional[torch.Tensor] = None,
                       out_ids: Optional[torch.Tensor] = None):
    """srag_exchange_merge_p2p: push this rank's exchange keys [B, k] into every rank's symmetric buffer over NVLink, wait
    for the others', merge -> (scores [B,k], ids [B,k]). peer_*_ptrs: ctypes uint64 arrays of `world` device addresses."""
    _need_cuda(local_keys)
    if local_keys.dtype != torch.int64 or local_keys.ndim != 2 or local_keys.shape[1] != k or not local_keys.is_contiguous():
        raise SragError("local_keys must be a contiguous int64 [B, k] tensor")
    B = local_keys.shape[0]
    dev = local_keys.device
    with torch.cuda.device(dev):
        os_ = out_scores if out_scores is not None else torch.empty(B, k, dtype=torch.float32, device=dev)
        oi = out_ids if out_ids is not None else torch.empty(B, k, dtype=torch.int64, device=dev)
        _lib.check(_lib.load().srag_exchange_merge_p2p(_ptr(local_keys), peer_key_ptrs, peer_flag_ptrs, world, rank, B, k,
                                                       epoch, cap_keys, cap_q, _ptr(os_), _ptr(oi), _stream(dev)),
                   "srag_exchange_merge_p2p")
    return os_, oi

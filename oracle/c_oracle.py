"""ctypes wrapper of the plain-C oracle (oracle/srag_oracle.c). TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "srag_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE] + (["-B"] if force else []), check=True, capture_output=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def scores(corpus, queries, metric="cosine"):
    lib = load()
    c = np.ascontiguousarray(corpus, dtype=np.float32)
    q = np.ascontiguousarray(queries, dtype=np.float32)
    out = np.empty((q.shape[0], c.shape[0]), dtype=np.float32)
    lib.oracle_scores_f32(_p(c), C.c_int64(c.shape[0]), _p(q), C.c_int(q.shape[0]), C.c_int(c.shape[1]),
                          C.c_int(0 if metric in (0, "cosine") else 1), _p(out))
    return out


def topk(score_rows, k, valid=None):
    lib = load()
    s = np.ascontiguousarray(score_rows, dtype=np.float32)
    B, n = s.shape
    out_s = np.full((B, k), -np.inf, dtype=np.float32)
    out_i = np.full((B, k), -1, dtype=np.int64)
    v = None if valid is None else np.ascontiguousarray(valid, dtype=np.uint8)
    lib.oracle_topk_row.restype = C.c_int
    for b in range(B):
        lib.oracle_topk_row(_p(s[b]), None if v is None else _p(v), C.c_int64(n), C.c_int(k), _p(out_s[b]), _p(out_i[b]))
    return out_s, out_i


def bm25_impacts(term_ptr, post_doc, post_tf, idf, dl, avgdl):
    lib = load()
    tp = np.ascontiguousarray(term_ptr, dtype=np.int64)
    pd = np.ascontiguousarray(post_doc, dtype=np.int32)
    tf = np.ascontiguousarray(post_tf, dtype=np.int32)
    idf = np.ascontiguousarray(idf, dtype=np.float32)
    dl = np.ascontiguousarray(dl, dtype=np.int32)
    imp = np.empty(pd.shape[0], dtype=np.float32)
    lib.oracle_bm25_impacts(_p(tp), C.c_int(tp.shape[0] - 1), _p(pd), _p(tf), _p(idf), _p(dl), C.c_float(avgdl), _p(imp))
    return imp


def bm25_scores(term_ptr, post_doc, post_imp, n_docs, terms):
    lib = load()
    tp = np.ascontiguousarray(term_ptr, dtype=np.int64)
    pd = np.ascontiguousarray(post_doc, dtype=np.int32)
    pi = np.ascontiguousarray(post_imp, dtype=np.float32)
    V = tp.shape[0] - 1
    t = np.array(sorted({int(x) for x in terms if 0 <= int(x) < V}), dtype=np.int32)
    acc = np.zeros(n_docs, dtype=np.float32)
    hit = np.zeros(n_docs, dtype=np.uint8)
    lib.oracle_bm25_accumulate(_p(tp), _p(pd), _p(pi), _p(t), C.c_int(t.shape[0]), _p(acc), _p(hit))
    return acc, hit.astype(bool)


def rrf(lists, rank_const=1, min_score=0.0):
    lib = load()
    width = max([len(l) for l in lists] + [1])
    ids = np.full((len(lists), width), -1, dtype=np.int64)
    for r, l in enumerate(lists):
        ids[r, :len(l)] = l
    out_i = np.empty(ids.size, dtype=np.int64)
    out_s = np.empty(ids.size, dtype=np.float64)
    lib.oracle_rrf.restype = C.c_int
    m = lib.oracle_rrf(_p(ids), C.c_int(len(lists)), C.c_int(width), C.c_int(rank_const), C.c_double(min_score),
                       _p(out_i), _p(out_s))
    return out_i[:m].copy(), out_s[:m].copy()

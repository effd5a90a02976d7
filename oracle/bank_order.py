"""CPU restatement of the index-build permutation `srag_bm25_bank_order` (csrc/bm25.cu: bm25_bank_order_kernel).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py). The reference has no BM25 index, so this follows
our own contract (include/srag_b200.h): inside every (term, sub-block) slice the postings are re-laid so that the 32
postings one warp of the scoring kernels touches per shared-memory access (positions 128g + 4l + e relative to the
slice start rounded down to a multiple of 4) fall on distinct banks (doc & 31) wherever the slice allows it.
The set of (doc, impact) pairs of a slice never changes, so scores cannot.
"""
from __future__ import annotations

import numpy as np


def round_robin_index(docs: np.ndarray) -> np.ndarray:
    """Position of every posting in the 'round-robin over banks' sequence: all first-of-their-bank postings in bank
    order, then all second ones, ... (rank inside a bank = slice order)."""
    n = len(docs)
    b = docs & 31
    c = np.bincount(b, minlength=32)
    run = np.zeros(32, dtype=np.int64)
    r = np.zeros(n, dtype=np.int64)
    for i in range(n):
        r[i] = run[b[i]]
        run[b[i]] += 1
    banks = np.arange(32)
    return np.array([np.minimum(c, r[i]).sum() + ((banks < b[i]) & (c > r[i])).sum() for i in range(n)], dtype=np.int64)


def access_position(s: int, lo_off: int, hi_off: int) -> int:
    """The s-th position of a slice occupying [lo_off, hi_off) (relative to its 4-aligned start) in ACCESS order:
    group of 128 first, then element e = pos & 3, then lane l = (pos & 127) >> 2."""
    g = (s + lo_off) >> 7
    i = s - ((128 * g - lo_off) if g else 0)
    sg = lo_off if g == 0 else 0
    tg = min(128, hi_off - 128 * g)
    for e in range(4):
        l0 = (sg - e + 3) >> 2
        l1 = (tg - e + 3) >> 2
        ne = max(0, l1 - l0)
        if i < ne:
            return 128 * g + 4 * (l0 + i) + e
        i -= ne
    raise AssertionError("position outside the slice")


def bank_order_slice(docs: np.ndarray, imps: np.ndarray, lo: int):
    """The slice [lo, lo + n) after the permutation: (docs, imps) in their new order."""
    n = len(docs)
    lo_off = lo & 3
    sidx = round_robin_index(np.asarray(docs, dtype=np.int64))
    out_d = np.empty_like(docs)
    out_w = np.empty_like(imps)
    for i in range(n):
        p = access_position(int(sidx[i]), lo_off, lo_off + n) - lo_off
        out_d[p] = docs[i]
        out_w[p] = imps[i]
    return out_d, out_w


def access_wavefronts(docs: np.ndarray, lo: int) -> tuple[int, int]:
    """(shared-memory wavefronts, accesses) of the accumulator accesses over one slice: an access = the postings at
    positions 128g + 4l + e for fixed (g, e); its cost = the largest number of them on one bank."""
    pos = np.arange(len(docs)) + (lo & 3)
    grp = (pos >> 7) * 4 + (pos & 3)
    tot = acc = 0
    for gkey in np.unique(grp):
        tot += int(np.bincount(np.asarray(docs)[grp == gkey] & 31, minlength=32).max())
        acc += 1
    return tot, acc

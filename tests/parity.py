"""Parity rules shared by the GPU tests (north_star): ids bit-exact wherever the score gap
exceeds the tolerance, cosine scores within 1e-3 absolute of the fp32 oracle."""
from __future__ import annotations

import numpy as np

TOL = 1e-3          # north_star tolerance
TIGHT = 5e-5        # our own bar: fp32 accumulation-order noise only


def check_topk_against_full_scores(gpu_scores, gpu_ids, full_scores, k, tol=TOL, id_base=0, valid=None):
    """gpu_scores/gpu_ids [B,k]; full_scores [B,N] oracle scores of EVERY row (fp32 or fp64).

    Asserts, per query:
      * every returned score is within tol of the oracle score of the returned id;
      * returned order is (score desc, id asc) in the GPU's own scores;
      * every row whose oracle score beats the oracle k-th by more than tol is returned;
      * no returned row is worse than the oracle k-th by more than tol;
      * position j holds exactly the oracle's id whenever its neighbours are > tol away.
    Returns the number of positions whose id differs from the strict oracle order.
    """
    gpu_scores = np.asarray(gpu_scores)
    gpu_ids = np.asarray(gpu_ids)
    B, N = full_scores.shape
    mismatches = 0
    for b in range(B):
        o = full_scores[b].astype(np.float64)
        if valid is not None:
            o = np.where(valid, o, -np.inf)
        kk = min(k, int(np.isfinite(o).sum()))
        order = np.lexsort((np.arange(N), -o))[:kk]
        got_ids = gpu_ids[b, :kk] - id_base
        got_s = gpu_scores[b, :kk].astype(np.float64)
        assert (gpu_ids[b, kk:] == -1).all(), f"q{b}: padding ids expected after {kk}"
        assert (got_ids >= 0).all() and (got_ids < N).all(), f"q{b}: id out of range {got_ids}"
        assert len(set(got_ids.tolist())) == kk, f"q{b}: duplicate ids {got_ids}"
        assert np.abs(got_s - o[got_ids]).max(initial=0) <= tol, \
            f"q{b}: score error {np.abs(got_s - o[got_ids]).max()}"
        keys = list(zip((-got_s).tolist(), got_ids.tolist()))
        assert keys == sorted(keys), f"q{b}: not sorted by (score desc, id asc)"
        if kk == 0:
            continue
        vk = o[order[-1]]
        must = set(np.nonzero(o > vk + tol)[0].tolist())
        assert must <= set(got_ids.tolist()), f"q{b}: missing clear winners {must - set(got_ids.tolist())}"
        assert (o[got_ids] >= vk - tol).all(), f"q{b}: returned a row more than tol below the k-th"
        so = o[order]
        for j in range(kk):
            if got_ids[j] != order[j]:
                mismatches += 1
                left = so[j - 1] - so[j] if j > 0 else np.inf
                right = so[j] - (so[j + 1] if j + 1 < kk else (np.sort(o)[::-1][kk] if N > kk else -np.inf))
                assert min(left, right) <= tol, \
                    f"q{b} pos {j}: id {got_ids[j]} != oracle {order[j]} although gaps are {left}, {right}"
    return mismatches


def full_scores_bf16(corpus_f32_rounded, queries_f32_rounded, metric="cosine"):
    """fp64 scores over bf16-rounded inputs (what both sides consume)."""
    import oracle
    return oracle.exact_topk_scores_f64(corpus_f32_rounded, queries_f32_rounded, metric)

"""K0 + K1 parity on the GPU, through the C ABI (super_rag_b200.ops -> libsrag_b200.so)."""
import os

import numpy as np
import pytest

import oracle
from parity import TIGHT, TOL, check_topk_against_full_scores

pytestmark = pytest.mark.gpu


def _run(C32, Q32, k, metric="cosine", cta_group=None, deleted=None, id_base=0):
    import torch
    from super_rag_b200 import ops
    if cta_group is None:
        os.environ.pop("SRAG_TOPK_CTA_GROUP", None)
    else:
        os.environ["SRAG_TOPK_CTA_GROUP"] = str(cta_group)
    dev = torch.device("cuda:0")
    c = torch.from_numpy(C32).to(dev)
    q = torch.from_numpy(Q32).to(dev)
    cb, cs = ops.ingest_rows(c, metric)
    qb, qs = ops.ingest_rows(q, metric)
    if deleted is not None and len(deleted):
        cs[torch.as_tensor(deleted, device=dev)] = float("nan")
    s, i = ops.topk_search(cb, cs, qb, qs, k, id_base=id_base)
    torch.cuda.synchronize()
    os.environ.pop("SRAG_TOPK_CTA_GROUP", None)
    return s.cpu().numpy(), i.cpu().numpy(), cb.float().cpu().numpy(), qb.float().cpu().numpy(), cs.cpu().numpy()


def _check(C32, Q32, k, metric="cosine", cta_group=None, deleted=None, id_base=0, tight=True):
    s, i, cr, qr, cs = _run(C32, Q32, k, metric, cta_group, deleted, id_base)
    assert (cr == oracle.bf16_round(C32)).all() and (qr == oracle.bf16_round(Q32)).all(), "bf16 cast differs"
    full = oracle.exact_topk_scores_f64(cr, qr, metric)
    valid = None
    if deleted is not None and len(deleted):
        valid = np.ones(C32.shape[0], bool)
        valid[np.asarray(deleted)] = False
    check_topk_against_full_scores(s, i, full, k, TOL, id_base, valid)
    if tight:
        check_topk_against_full_scores(s, i, full, k, TIGHT, id_base, valid)
    # the fp32 oracle itself (blocked BLAS) agrees under the same rule
    os_, oi = oracle.exact_topk(cr, qr, k, metric, id_base=id_base, valid=valid)
    same = (oi == i).mean()
    assert same > 0.98, f"only {same:.3f} of ids equal the fp32 oracle's"
    return s, i


def _data(n, d, b, seed, planted=True):
    rng = np.random.default_rng(seed)
    C = rng.standard_normal((n, d)).astype(np.float32)
    Q = rng.standard_normal((b, d)).astype(np.float32)
    if planted and n > 0:
        h = b // 2
        src = rng.integers(0, n, size=h)
        Q[:h] = C[src] + 0.1 * rng.standard_normal((h, d)).astype(np.float32)
    return C, Q


def test_ingest_scale_matches_oracle():
    import torch
    from super_rag_b200 import ops
    C, _ = _data(1000, 256, 4, 0)
    C[3] = 0
    c = torch.from_numpy(C).cuda()
    cb, cs = ops.ingest_rows(c, "cosine")
    cr = cb.float().cpu().numpy()
    assert (cr == oracle.bf16_round(C)).all()
    want = oracle.row_scales(cr, "cosine")
    got = cs.cpu().numpy()
    assert got[3] == 0.0
    assert np.allclose(got, want, rtol=2e-6, atol=0)
    _, cs2 = ops.ingest_rows(cb, "cosine")           # bf16 source path
    assert np.allclose(cs2.cpu().numpy(), want, rtol=2e-6, atol=0)
    _, ci = ops.ingest_rows(c, "ip")
    assert (ci.cpu().numpy() == 1.0).all()


@pytest.mark.parametrize("cta_group", [1, 2])
def test_small_edge_cases(cta_group):
    C, Q = _data(3000, 64, 8, 1)
    C[17] = 0.0                 # zero-norm row scores 0
    C[2000] = C[5]              # exact duplicate -> tie broken by id
    Q[0] = C[5] * 2.0
    Q[7] = 0.0                  # zero query: all scores 0, ids ascending
    s, i = _check(C, Q, 7, cta_group=cta_group)
    assert i[0, 0] == 5 and i[0, 1] == 2000
    assert (s[7] == 0).all() and i[7].tolist() == list(range(7))


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("n,d,b,k", [(5000, 1024, 100, 10), (777, 200, 3, 5), (256, 64, 1, 1),
                                      (40000, 1024, 130, 100), (1000, 128, 260, 128)])
def test_parity_shapes(n, d, b, k, cta_group):
    C, Q = _data(n, d, b, n + d + b)
    _check(C, Q, k, cta_group=cta_group)


def test_auto_group_large_batch():
    C, Q = _data(30000, 1024, 1024, 5)
    _check(C, Q, 10)


def test_inner_product_and_id_base():
    C, Q = _data(4000, 256, 20, 9)
    _check(C, Q, 10, metric="ip", id_base=1_000_000_000_000, tight=False)


def test_tombstones_and_short_corpus():
    C, Q = _data(600, 64, 5, 11)
    deleted = [0, 5, 17, 599]
    Q[0] = C[5]
    s, i = _check(C, Q, 10, deleted=deleted)
    assert not set(i.reshape(-1).tolist()) & set(deleted)
    C2, Q2 = _data(6, 64, 3, 12)
    s, i = _check(C2, Q2, 10)       # N < k -> padded with (-inf, -1)
    assert (i[:, 6:] == -1).all() and np.isinf(s[:, 6:]).all()


def test_many_duplicates_tie_by_id():
    rng = np.random.default_rng(3)
    base = rng.standard_normal((50, 128)).astype(np.float32)
    C = np.tile(base, (40, 1))                     # every row appears 40 times
    Q = base[:10].copy()
    s, i = _check(C, Q, 64)
    for b in range(10):
        assert i[b, :40].tolist() == [b + 50 * j for j in range(40)]


def test_bad_arguments_raise():
    import torch
    from super_rag_b200 import ops, SragError
    c = torch.zeros(10, 64, device="cuda", dtype=torch.bfloat16)
    s = torch.ones(10, device="cuda")
    with pytest.raises(SragError):
        ops.topk_search(c, s, c[:2], s[:2], 129)
    with pytest.raises(SragError):
        ops.topk_search(c.cpu(), s, c[:2], s[:2], 5)
    with pytest.raises(SragError):
        ops.topk_search(c, s, torch.zeros(2, 32, device="cuda", dtype=torch.bfloat16), s[:2], 5)


@pytest.mark.parametrize("n,d,b,k", [
    (3000, 8, 5, 3),            # smallest legal dimension
    (3000, 72, 9, 4),           # dim not a multiple of the 64-wide k-block: TMA zero-fills the tail
    (900, 4096, 4, 6),          # 64 k-blocks
    (2000, 64, 129, 5),         # ragged batches: zero-padded query staging, one and two q-tiles
    (2000, 64, 257, 5),
    (50, 64, 1030, 3),          # more queries than one call takes: ops splits the batch
])
def test_parity_more_shapes(n, d, b, k):
    C, Q = _data(n, d, b, 7 * n + d + b)
    _check(C, Q, k)


def test_single_q_tile_large_k_uses_shared_threshold():
    """B <= 128 with k >= 32: 148 lists per query publish their best score (threshold sharing across all
    slabs). Planted duplicates and a cluster of near neighbours make sure nothing above the shared bound is lost."""
    C, Q = _data(300_000, 128, 100, 99)
    Q[:50] = C[1000:1050] + 0.05 * Q[:50]
    C[250_000:250_050] = C[1000:1050]          # exact duplicates far away: both must be returned, lower id first
    s, i = _check(C, Q, 100)
    for b in range(50):
        assert i[b, 0] == 1000 + b and i[b, 1] == 250_000 + b


def test_empty_corpus_returns_padding():
    import torch
    from super_rag_b200 import ops
    dev = torch.device("cuda:0")
    q = torch.randn(3, 64, device=dev)
    qb, qs = ops.ingest_rows(q, "cosine")
    s, i = ops.topk_search(torch.empty(0, 64, dtype=torch.bfloat16, device=dev), torch.empty(0, device=dev), qb, qs, 4)
    assert (i.cpu().numpy() == -1).all() and np.isinf(s.cpu().numpy()).all()


def test_k1_against_reference_ranked_lists(topk_golden):
    """K1 against ranked lists produced by the REFERENCE's own brute-force cosine ranking (tests/golden/topk_golden.json:
    search_utils.py:337-366 with calculate_cosine_similarity :56-67, executed by tests/golden/make_golden.py). The inputs are
    exact in bf16, so the kernel consumes the very numbers the reference function saw: scores within 1e-3 (north_star;
    5e-5 in fact), ids equal position by position wherever the reference's scores separate neighbours by more than the
    tolerance, duplicates in row order, zero rows and the zero query score 0."""
    C, cases, k = topk_golden
    Q = np.stack([c["query"] for c in cases])
    ref_full = np.stack([c["scores"] for c in cases])
    for cta_group in (None, 1):
        s, i, cr, qr, _ = _run(C, Q, k, "cosine", cta_group)
        assert (cr == C).all() and (qr == Q).all()
        check_topk_against_full_scores(s, i, ref_full, k, TOL)
        check_topk_against_full_scores(s, i, ref_full, k, TIGHT)
        assert i[0, :2].tolist() == [5, 1200] and i[1, :3].tolist() == [300, 1500, 1900]     # duplicates: lower row first
        assert i[3].tolist() == list(range(k)) and (s[3] == 0).all()                          # zero query: row order
        exact = sum(int((i[qi] == c["top_ids"]).all()) for qi, c in enumerate(cases) if c["min_score"] < -1.5)
        assert exact >= 8, exact
        for qi, c in enumerate(cases):                    # the branch's `score > min_score` cut on our ranked list
            if c["min_score"] > -1.5:
                kept = [int(r) for r, x in zip(i[qi], s[qi]) if x > c["min_score"]]
                assert kept == c["top_ids"].tolist()[:len(kept)] and abs(len(kept) - len(c["top_ids"])) <= 1
    # through the batched connector API as well, and a single-row batch (B = 1 plan)
    import torch
    from super_rag_b200.vectorstore import B200Client
    cl = B200Client()
    cl.delete_collection("t_golden")
    col = cl.create_collection("t_golden", C.shape[1], "cosine")
    col.append(torch.from_numpy(C))
    for qi in (0, 1, 3, 5):
        s1, i1 = col.search_batch(torch.from_numpy(Q[qi:qi + 1]), k)
        check_topk_against_full_scores(s1.cpu().numpy(), i1.cpu().numpy(), ref_full[qi:qi + 1], k, TIGHT)
    cl.delete_collection("t_golden")


@pytest.mark.parametrize("b,k", [(1024, 10), (1024, 100), (64, 10), (64, 100), (300, 33)])
def test_shared_bound_ties_at_the_boundary_and_skewed_lists(b, k):
    """The shared bound is the ceil(k/j)-th largest of the lists' published j-th best scores and tile 0 is scanned twice
    (maxima first). Adversarial layouts for it: (a) hundreds of EXACT copies of the query's best row spread over every
    list, so the k-th boundary falls inside one tie group that spans all lists (lowest rows must win); (b) all of a
    query's top-k inside ONE tile of one list while every other list publishes low maxima; (c) copies inside tile 0 of a
    list only (what the prescan sees) with better rows arriving later."""
    rng = np.random.default_rng(1000 * b + k)
    n, d = 120_000, 64
    C = rng.standard_normal((n, d)).astype(np.float32)
    Q = rng.standard_normal((b, d)).astype(np.float32)
    v = rng.standard_normal((3, d)).astype(np.float32)
    spread = np.arange(37, n, 301)[:390]                 # (a) 390 copies, one every 301 rows: every tile range, every list
    C[spread] = v[0]
    Q[0] = v[0]
    C[70_000:70_000 + 2 * k] = v[1] + 0.01 * rng.standard_normal((2 * k, d)).astype(np.float32)   # (b) one tile of one list
    Q[1] = v[1]
    C[1000:1000 + k // 2 + 1] = v[2] + 0.2 * rng.standard_normal((k // 2 + 1, d)).astype(np.float32)  # (c) weak copies in a first tile ...
    C[100_000:100_000 + k] = v[2] + 0.01 * rng.standard_normal((k, d)).astype(np.float32)          # ... better ones much later
    Q[2] = v[2]
    Q[b - 1] = v[0]                                      # the last lane of the last q-tile sees the tie group too
    s, i = _check(C, Q, k)
    assert i[0].tolist() == spread[:k].tolist() and i[b - 1].tolist() == spread[:k].tolist()
    assert set(i[1].tolist()) <= set(range(70_000, 70_000 + 2 * k))
    assert set(i[2].tolist()) == set(range(100_000, 100_000 + k))


@pytest.mark.parametrize("env", [
    {"SRAG_TOPK_RQ": "1"},                                   # resident-query plan (queries parked in TMEM + shared memory)
    {"SRAG_TOPK_EPI_WARPS": "8"},                            # two column slices per unit: twice the lists per query
    {"SRAG_TOPK_RQ": "1", "SRAG_TOPK_EPI_WARPS": "8"},
    {"SRAG_TOPK_PRESCAN": "0"},                              # no tile-0 prescan: the bound appears after the first tiles
    {"SRAG_TOPK_SHARE": "0"},                                # every list on its own k-th best
    {"SRAG_TOPK_SHARE": "2", "SRAG_TOPK_SHARE_PERIOD": "1"}, # sharing wherever possible, bound re-read on every tile
    {"SRAG_TOPK_WARM": "4"},                                 # warm-up tiles (k >= 32 instantiation) on top of the prescan
    {"SRAG_TOPK_WINDOW": "0"},                               # no L2 pacing
], ids=lambda e: ",".join(f"{k[10:]}={v}" for k, v in e.items()))
def test_selectable_plans_and_knobs_keep_parity(env):
    """Every plan / knob that DESIGN.md calls selectable returns the same exact top-k: both unit shapes, top-10 and
    top-100, ties and tombstones included (the knobs are read per call, so they can be set in-process)."""
    rng = np.random.default_rng(77)
    n, d = 30_000, 128
    C = rng.standard_normal((n, d)).astype(np.float32)
    C[15_000:15_040] = C[5:45]                       # exact duplicates: lower row first
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        for b, k in [(300, 10), (1024, 100), (64, 100), (9, 1)]:
            Q = rng.standard_normal((b, d)).astype(np.float32)
            Q[:5] = C[5:10] + 0.01 * Q[:5]
            s, i = _check(C, Q, k, deleted=[7, 15_002])      # row 7 and its copy: query 2's two exact matches are gone
            if k >= 2:
                assert i[0, 0] == 5 and i[0, 1] == 15_000 and i[1, 0] == 6 and i[1, 1] == 15_001
                assert 7 not in i[2] and 15_002 not in i[2]
    finally:
        for k_, v in old.items():
            if v is None:
                os.environ.pop(k_, None)
            else:
                os.environ[k_] = v

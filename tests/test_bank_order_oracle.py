"""CPU checks of the bank-order restatement (oracle/bank_order.py): the access-order map is a bijection, the
permutation keeps every slice's set of postings, and it cuts the modelled shared-memory wavefronts."""
import numpy as np

from oracle import bank_order as bo


def test_access_position_is_a_bijection_in_access_order():
    for lo_off in range(4):
        for n in list(range(1, 140)) + [255, 256, 257, 511, 700]:
            hi_off = lo_off + n
            ps = [bo.access_position(s, lo_off, hi_off) for s in range(n)]
            assert sorted(ps) == list(range(lo_off, hi_off))
            keys = [((p >> 7), p & 3, (p & 127) >> 2) for p in ps]
            assert keys == sorted(keys)


def test_round_robin_sequence_puts_distinct_banks_first():
    rng = np.random.default_rng(0)
    docs = np.sort(rng.choice(2048, 300, replace=False))
    s = bo.round_robin_index(docs)
    assert sorted(s.tolist()) == list(range(300))
    order = np.argsort(s)
    banks = docs[order] & 31
    full_rows = np.bincount(docs & 31, minlength=32).min()
    for r in range(full_rows):
        assert sorted(banks[32 * r:32 * r + 32].tolist()) == list(range(32))


def test_bank_order_keeps_the_slice_and_cuts_wavefronts():
    rng = np.random.default_rng(1)
    before = after = 0
    for n, lo in [(20, 0), (100, 3), (513, 1), (840, 2), (2048, 0)]:
        docs = np.sort(rng.choice(2048, n, replace=False)).astype(np.int32) + 4096
        imps = rng.random(n).astype(np.float32)
        d2, w2 = bo.bank_order_slice(docs, imps, lo)
        o = np.argsort(d2, kind="stable")
        assert (d2[o] == docs).all() and (w2[o] == imps).all()
        before += bo.access_wavefronts(docs, lo)[0]
        after += bo.access_wavefronts(d2, lo)[0]
    assert after < 0.6 * before, (before, after)
    full = np.arange(2048, dtype=np.int32)          # every document present: perfectly conflict-free afterwards
    d2, _ = bo.bank_order_slice(full, np.zeros(2048, np.float32), 0)
    w, a = bo.access_wavefronts(d2, 0)
    assert w == a

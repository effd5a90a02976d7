"""srag_topk_search_static: the exact top-k scan with the library's own static device scratch, for callers that may not
allocate device memory (CUDA graph capture, tight serving loops). No counterpart in the reference
(include/srag_b200.h). Collected last on purpose.

Called through the C ABI (ctypes) with caller-owned inputs and outputs, twice in a row on the shared scratch — the second
call smaller than the first, so stale heaps of the first would show — and compared with the existing entry point
(srag_topk_search with a caller workspace): ids identical, scores within the suite's own tolerance (parity.TIGHT; they
are in fact bit-equal). Also: sizes beyond the static array are rejected with a clear error, the driver's free-memory
figure does not move across a call, and the call can be captured into a CUDA graph and replayed.
"""
import ctypes as C

import numpy as np
import pytest

from parity import TIGHT

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _setup(n, d, seed):
    import torch
    from super_rag_b200 import ops
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    corpus = torch.randn(n, d, generator=g, device="cuda")
    corpus[n // 2] = corpus[3]                      # an exact duplicate: the lower row must come first in both paths
    cb, cs = ops.ingest_rows(corpus, "cosine")
    cs[7] = float("nan")                            # a tombstone
    return corpus, cb, cs, g


def _queries(corpus, b, g):
    import torch
    from super_rag_b200 import ops
    q = torch.randn(b, corpus.shape[1], generator=g, device="cuda")
    h = b // 2
    src = torch.randint(0, corpus.shape[0], (h,), generator=g, device="cuda")
    q[:h] = corpus[src] + 0.1 * q[:h]
    q[0] = corpus[3]
    return ops.ingest_rows(q, "cosine")


def _static(lib, cb, cs, qb, qs, k, out_s, out_i, id_base, stream):
    return lib.srag_topk_search_static(_ptr(cb), _ptr(cs), cb.shape[0], cb.shape[1], _ptr(qb), _ptr(qs), qb.shape[0], k,
                                       id_base, _ptr(out_s), _ptr(out_i), C.c_void_p(stream))


def test_static_scratch_search_matches_workspace_search_twice_in_a_row():
    import torch
    from super_rag_b200 import _lib, ops
    lib = _lib.load()
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    n, d = 150_003, 256
    corpus, cb, cs, g = _setup(n, d, seed=91)
    # call 1: four query tiles (CTA pairs), k = 10; call 2: a ragged single tile, k = 5 — smaller, on the same scratch
    for b, k, id_base in ((700, 10, 0), (37, 5, 1_000_000)):
        qb, qs = _queries(corpus, b, g)
        want_s, want_i = ops.topk_search(cb, cs, qb, qs, k, id_base=id_base)          # existing entry point
        out_s = torch.full((b, k), 123.0, dtype=torch.float32, device="cuda")
        out_i = torch.full((b, k), -7, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        rc = _static(lib, cb, cs, qb, qs, k, out_s, out_i, id_base, stream)
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert rc == 0, lib.srag_last_error()
        assert torch.equal(out_i, want_i), (b, k)
        assert (out_s - want_s).abs().max().item() <= TIGHT
        assert torch.equal(out_s.view(torch.int32), want_s.view(torch.int32))         # same kernel, same arithmetic
        assert out_i[0, 0].item() == id_base + 3 and out_i[0, 1].item() == id_base + n // 2
        assert not (out_i == id_base + 7).any()                                       # the tombstone never comes back
        if (b, k) != (700, 10):          # (the very first call may load the module lazily)
            assert free_after == free_before, "the call changed the device's free memory"
    # the ops wrapper allocates nothing either and gives the same answer
    out_s2, out_i2 = torch.empty_like(out_s), torch.empty_like(out_i)
    ops.topk_search_static(cb, cs, qb, qs, 5, out_s2, out_i2, id_base=1_000_000)
    torch.cuda.synchronize()
    assert torch.equal(out_i2, out_i) and torch.equal(out_s2, out_s)


def test_static_scratch_search_rejects_what_does_not_fit_and_bad_arguments():
    import torch
    from super_rag_b200 import _lib, ops
    lib = _lib.load()
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    corpus, cb, cs, g = _setup(20_000, 128, seed=5)
    qb, qs = _queries(corpus, 1024, g)
    k = 100
    need = ops.topk_workspace_bytes(cb.shape[0], 1024, k)
    assert need > ops.TOPK_STATIC_SCRATCH_BYTES                  # 1024 queries x top-100 needs ~32 MB of heaps
    out_s = torch.empty(1024, k, dtype=torch.float32, device="cuda")
    out_i = torch.empty(1024, k, dtype=torch.int64, device="cuda")
    rc = _static(lib, cb, cs, qb, qs, k, out_s, out_i, 0, stream)
    msg = lib.srag_last_error().decode()
    assert rc == -3 and "SRAG_TOPK_STATIC_SCRATCH_BYTES" in msg and str(need) in msg, (rc, msg)
    with pytest.raises(ops.SragError, match="static array"):
        ops.topk_search_static(cb, cs, qb, qs, k, out_s, out_i)
    assert lib.srag_topk_search_static(_ptr(cb), _ptr(cs), cb.shape[0], cb.shape[1], _ptr(qb), _ptr(qs), 1024, 10, 0,
                                       C.c_void_p(0), _ptr(out_i), C.c_void_p(stream)) == -1      # null output
    # what fits the documented limit does fit: 1024 queries at k = 10 over any corpus size
    assert ops.topk_workspace_bytes(100_000_000, 1024, 10) <= ops.TOPK_STATIC_SCRATCH_BYTES
    torch.cuda.synchronize()


def test_static_scratch_search_under_cuda_graph_capture():
    """The use case: nothing in the call allocates, so it can be captured and replayed (with new queries written into the
    captured buffers)."""
    import torch
    from super_rag_b200 import _lib, ops
    lib = _lib.load()
    torch.cuda.set_device(0)
    corpus, cb, cs, g = _setup(60_000, 256, seed=17)
    b, k = 200, 10
    qb, qs = _queries(corpus, b, g)
    out_s = torch.empty(b, k, dtype=torch.float32, device="cuda")
    out_i = torch.empty(b, k, dtype=torch.int64, device="cuda")
    ops.topk_search_static(cb, cs, qb, qs, k, out_s, out_i)      # set-up outside capture (symbol address, smem opt-in)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        rc = _static(lib, cb, cs, qb, qs, k, out_s, out_i, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.srag_last_error()
    for _ in range(2):
        qb2, qs2 = _queries(corpus, b, g)
        qb.copy_(qb2)
        qs.copy_(qs2)
        out_i.fill_(-9)
        graph.replay()
        torch.cuda.synchronize()
        want_s, want_i = ops.topk_search(cb, cs, qb, qs, k)
        assert torch.equal(out_i, want_i) and torch.equal(out_s.view(torch.int32), want_s.view(torch.int32))
    del graph

#!/usr/bin/env python
"""A/B timing of K2 builds on ONE index: the C3 index is built once with the default library, then every library given on
the command line (tools/build_variant.sh outputs) runs srag_bm25_search on the same device arrays; results must be bit-equal
to the default library's.

    python tools/ab_k2.py [--docs N] [--stop] lib_a.so lib_b.so ...
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_hybrid as bh  # noqa: E402
from super_rag_b200 import _lib, ops  # noqa: E402


def bind(path):
    lib = C.CDLL(path)
    for name in ("srag_bm25_search", "srag_bm25_workspace_bytes", "srag_last_error"):
        fn = getattr(lib, name)
        fn.argtypes = _lib.SIGNATURES[name]
        fn.restype = _lib._RESTYPE.get(name, C.c_int)
    return lib


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--stop", action="store_true", help="also time the stop-word queries")
    ap.add_argument("libs", nargs="*")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    idx = bh.build_index(args.docs, dev)
    sets = [("c3", bh.make_queries(1024, 8, dev))]
    if args.stop:
        sets.append(("stop", bh.make_queries(1024, 8, dev, lo_rank=1, seed=5)))
    k = 100
    ws = torch.empty(ops.bm25_workspace_bytes(args.docs, 1024, k), dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for tag, qt in sets:
        post_bytes = idx.query_posting_bytes(qt)
        ref = None
        for path in [_lib.lib_path()] + args.libs:
            lib = bind(path)
            s = torch.empty(1024, k, dtype=torch.float32, device=dev)
            i = torch.empty(1024, k, dtype=torch.int64, device=dev)

            def run():
                rc = lib.srag_bm25_search(idx.slice_ptr.data_ptr(), idx.post_off.data_ptr(), idx.post_imp.data_ptr(), idx.vocab,
                                          idx.n_docs, qt.data_ptr(), 1024, qt.shape[1], k, 0, s.data_ptr(), i.data_ptr(),
                                          ws.data_ptr(), ws.numel(), st)
                assert rc == 0, lib.srag_last_error()
            ms, _ = bh.timed(run, 2, args.steps)
            same = None
            if ref is None:
                ref = (s.clone(), i.clone())
            else:
                same = bool(torch.equal(ref[1], i)) and bool(torch.equal(ref[0].view(torch.int32), s.view(torch.int32)))
            print(json.dumps({"set": tag, "lib": os.path.basename(path), "ms": round(ms, 3),
                              "algorithmic_gbs": round(post_bytes / ms / 1e6, 1), "same_as_default": same}), flush=True)


if __name__ == "__main__":
    main()

"""ctypes binding of libsrag_b200.so (include/srag_b200.h). There is NO CPU fallback: if the
library cannot be loaded, or a call fails, a ``SragError`` is raised."""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None


class SragError(RuntimeError):
    """A libsrag_b200 call returned a negative status (message from srag_last_error())."""


_i64, _i32, _f32, _f64, _vp, _sz = C.c_int64, C.c_int, C.c_float, C.c_double, C.c_void_p, C.c_size_t

# name -> (argtypes). restype is int unless listed in _RESTYPE
SIGNATURES = {
    "srag_abi_version": [],
    "srag_last_error": [],
    "srag_device_info": [C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32)],
    "srag_rows_ingest_f32": [_vp, _i64, _i32, _vp, _vp, _i32, _vp],
    "srag_rows_scale_bf16": [_vp, _i64, _i32, _vp, _i32, _vp],
    "srag_topk_workspace_bytes": [_i64, _i32, _i32, C.POINTER(_sz)],
    "srag_topk_workspace_bytes_ex": [_i64, _i32, _i32, _i32, C.POINTER(_sz)],
    "srag_topk_search": [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _i32, _i64, _vp, _vp, _vp, _sz, _vp],
    "srag_topk_search_keys": [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _i32, _i64, _vp, _vp, _vp, _vp, _sz, _vp],
    "srag_topk_search_static": [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _i32, _i64, _vp, _vp, _vp],
    "srag_topk_search_filtered": [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                  _sz, _vp],
    "srag_mmr_rerank": [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _vp, _i32, _i64, _f32, _f32, _vp, _vp, _vp],
    "srag_prof_enable": [_i32],
    "srag_prof_read": [C.POINTER(_f64), C.POINTER(_f64), C.POINTER(_i32)],
    "srag_bm25_impacts": [_vp, _i32, _vp, _vp, _i64, _vp, _vp, _f32, _vp, _vp],
    "srag_bm25_block_index": [_vp, _vp, _i32, _i64, _i32, _vp, _vp],
    "srag_bm25_block_docs": [],
    "srag_bm25_pack_plan": [_vp, _i32, _i64, _vp, _vp, _vp],
    "srag_bm25_pack_fill": [_vp, _vp, _i32, _i64, _vp, _vp, _vp, _vp, _i32, _vp],
    "srag_bm25_unpack_docs": [_vp, _i32, _i64, _vp, _vp, _vp, _vp, _vp],
    "srag_bm25_workspace_bytes": [_i64, _i32, _i32, C.POINTER(_sz)],
    "srag_bm25_build_plan": [_vp, _vp, _i64, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "srag_bm25_build_fill": [_vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp, _vp, _vp, _vp],
    "srag_bm25_search": [_vp, _vp, _vp, _i32, _i64, _vp, _i32, _i32, _i32, _i64, _vp, _vp, _vp, _sz, _vp],
    "srag_bm25_search_filtered": [_vp, _vp, _vp, _i32, _i64, _vp, _i32, _i32, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp,
                                  _vp, _sz, _vp],
    "srag_fuse": [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _f64, C.POINTER(_f64), _i32, _vp, _vp, _vp],
    "srag_merge_candidates": [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp],
    "srag_merge_keys": [_vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp],
    "srag_pack_keys": [_vp, _vp, _i64, _vp, _vp, _vp],
    "srag_exchange_merge_p2p": [_vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_uint32, _sz, _i32, _vp, _vp, _vp],
}
_RESTYPE = {"srag_last_error": C.c_char_p}


def lib_path() -> str:
    return _build.LIB_PATH


def load():
    """Load (building first if the in-tree .so is missing or stale and nvcc is present)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = _build.LIB_PATH
        override = os.environ.get("SRAG_B200_LIB")  # A/B experiments: load exactly this build
        if override:
            path = override
        elif not _build.is_fresh():
            try:
                path = _build.build()
            except Exception as e:  # no nvcc on this machine: use the shipped .so if there is one
                if not os.path.exists(path):
                    raise SragError(
                        f"libsrag_b200.so is missing and could not be built ({e}); the B200 kernels are the "
                        "only implementation — there is no CPU fallback") from e
        lib = C.CDLL(path)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the header and the library disagree
            fn.argtypes = argtypes
            fn.restype = _RESTYPE.get(name, C.c_int)
        if lib.srag_abi_version() != 2:
            raise SragError(f"ABI version mismatch: library reports {lib.srag_abi_version()}")
        _lib = lib
    return _lib


def check(status: int, what: str):
    if status != 0:
        msg = load().srag_last_error()
        raise SragError(f"{what} failed ({status}): {msg.decode() if msg else ''}")

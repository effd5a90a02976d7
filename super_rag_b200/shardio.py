"""On-disk shard format: the device-resident index of one rank, written so that a
serving process can restart without re-embedding or re-tokenising (checkpoint / resume).

The reference keeps this state inside the SeekDB server (`collection.add`, seekdb_connector.py:68-85); this is
the stand-alone equivalent for the B200 store. One file per shard:

    [ "SRAGSHD1" ][ section 0 | pad to 4096 ][ section 1 | pad ] ... [ footer JSON ][ u64 footer length ][ "SRAGSHD1" ]

Sections are raw little-endian arrays (bf16 is stored as its 16-bit pattern) or JSON blobs; the footer lists
`{name, dtype, shape, offset, nbytes, crc32}` per section plus a free-form `meta` dict. Every section carries a
CRC-32 that `read_shard` verifies, the file is written to `<path>.tmp` and renamed, so a crash never leaves a
half-written shard under the final name. Arrays stream between the device and the file in bounded chunks
through one pinned staging buffer, so saving a 20 GB corpus does not need 20 GB of host memory.

Kinds written by this package:
  "vector"  rows bf16 [n, dim], row_scale fp32 [n] (NaN = tombstone), doc table (ids / texts / metadata)
  "bm25"    the searched layout as it sits in HBM: term_ptr int64 [V+1], slice_ptr uint32 [V, n_sub+1], post_off uint16,
            post_imp fp32 (padded slices), dl int32 [n_docs], df int64 [V]; round-1 shards (post_doc int32) still load
"""
from __future__ import annotations

import json
import os
import struct
import zlib
from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch

MAGIC = b"SRAGSHD1"
VERSION = 1
ALIGN = 4096
CHUNK_BYTES = 64 << 20

_TORCH_TO_NAME = {
    torch.bfloat16: "bf16", torch.float32: "f32", torch.float64: "f64", torch.int32: "i32", torch.int64: "i64",
    torch.uint8: "u8", torch.int16: "i16", torch.float16: "f16",
}
_NAME_TO_TORCH = {v: k for k, v in _TORCH_TO_NAME.items()}
_ITEMSIZE = {"bf16": 2, "f16": 2, "f32": 4, "f64": 8, "i32": 4, "i64": 8, "u8": 1, "i16": 2}


class ShardError(RuntimeError):
    pass


def _as_bytes_view(t: torch.Tensor) -> torch.Tensor:
    """1-D uint8 view of a contiguous tensor (any dtype, any device)."""
    return t.contiguous().view(-1).view(torch.uint8)


def _pad(f, align=ALIGN):
    pos = f.tell()
    rem = (-pos) % align
    if rem:
        f.write(b"\0" * rem)


def write_shard(path: str, kind: str, meta: Dict[str, Any], arrays: Dict[str, torch.Tensor],
                blobs: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """Write one shard atomically. `arrays`: tensors on any device; `blobs`: JSON-serialisable objects.
    Returns the footer that was written."""
    sections = []
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC)
        for name, t in arrays.items():
            if not isinstance(t, torch.Tensor):
                t = torch.as_tensor(t)
            if t.dtype not in _TORCH_TO_NAME:
                raise ShardError(f"section {name!r}: unsupported dtype {t.dtype}")
            _pad(f)
            off = f.tell()
            raw = _as_bytes_view(t)
            crc = 0
            for a in range(0, raw.numel(), CHUNK_BYTES):
                piece = raw[a:a + CHUNK_BYTES].cpu().numpy()      # device -> host, one bounded chunk at a time
                crc = zlib.crc32(piece, crc)
                f.write(piece.data)
            sections.append({"name": name, "dtype": _TORCH_TO_NAME[t.dtype], "shape": list(t.shape), "offset": off,
                             "nbytes": int(raw.numel()), "crc32": crc & 0xFFFFFFFF})
        for name, obj in (blobs or {}).items():
            _pad(f)
            off = f.tell()
            data = json.dumps(obj, ensure_ascii=False, separators=(",", ":")).encode("utf-8")
            f.write(data)
            sections.append({"name": name, "dtype": "json", "shape": [len(data)], "offset": off, "nbytes": len(data),
                             "crc32": zlib.crc32(data) & 0xFFFFFFFF})
        _pad(f)
        footer = {"format": "srag_b200.shard", "version": VERSION, "kind": kind, "meta": meta, "sections": sections}
        data = json.dumps(footer, separators=(",", ":")).encode("utf-8")
        f.write(data)
        f.write(struct.pack("<Q", len(data)))
        f.write(MAGIC)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)
    return footer


def read_footer(path: str) -> Dict[str, Any]:
    size = os.path.getsize(path)
    if size < 2 * len(MAGIC) + 8:
        raise ShardError(f"{path}: too short to be a shard")
    with open(path, "rb") as f:
        if f.read(len(MAGIC)) != MAGIC:
            raise ShardError(f"{path}: bad magic")
        f.seek(size - len(MAGIC) - 8)
        (flen,) = struct.unpack("<Q", f.read(8))
        if f.read(len(MAGIC)) != MAGIC:
            raise ShardError(f"{path}: truncated (no trailer)")
        if flen > size - 2 * len(MAGIC) - 8:
            raise ShardError(f"{path}: corrupt footer length")
        f.seek(size - len(MAGIC) - 8 - flen)
        footer = json.loads(f.read(flen).decode("utf-8"))
    if footer.get("format") != "srag_b200.shard" or footer.get("version", 0) > VERSION:
        raise ShardError(f"{path}: unknown format / version {footer.get('version')}")
    return footer


def read_shard(path: str, device: Optional[torch.device] = None, expect_kind: Optional[str] = None,
               verify: bool = True, host_sections=None) -> Tuple[Dict[str, Any], Dict[str, torch.Tensor], Dict[str, Any]]:
    """-> (meta, arrays on `device` (host if None), blobs). Raises ShardError on any CRC / layout mismatch.
    `host_sections`: predicate on the section name; matching arrays stay in host memory (the doc-table columns)."""
    footer = read_footer(path)
    if expect_kind is not None and footer["kind"] != expect_kind:
        raise ShardError(f"{path}: kind {footer['kind']!r}, expected {expect_kind!r}")
    arrays: Dict[str, torch.Tensor] = {}
    blobs: Dict[str, Any] = {}
    dev = torch.device(device) if device is not None else None
    on_gpu = dev is not None and dev.type == "cuda"
    stage = torch.empty(CHUNK_BYTES, dtype=torch.uint8, pin_memory=True) if on_gpu else None
    with open(path, "rb") as f:
        for s in footer["sections"]:
            f.seek(s["offset"])
            if s["dtype"] == "json":
                data = f.read(s["nbytes"])
                if len(data) != s["nbytes"] or (verify and (zlib.crc32(data) & 0xFFFFFFFF) != s["crc32"]):
                    raise ShardError(f"{path}: section {s['name']!r} is corrupt")
                blobs[s["name"]] = json.loads(data.decode("utf-8"))
                continue
            dt = _NAME_TO_TORCH.get(s["dtype"])
            n_el = int(np.prod(s["shape"])) if s["shape"] else 1
            if dt is None or n_el * _ITEMSIZE[s["dtype"]] != s["nbytes"]:
                raise ShardError(f"{path}: section {s['name']!r} has an inconsistent header")
            to_gpu = on_gpu and not (host_sections is not None and host_sections(s["name"]))
            out = torch.empty(s["nbytes"], dtype=torch.uint8, device=dev if to_gpu else "cpu")
            crc = 0
            for a in range(0, s["nbytes"], CHUNK_BYTES):
                m = min(CHUNK_BYTES, s["nbytes"] - a)
                if to_gpu:
                    torch.cuda.current_stream(dev).synchronize()        # staging buffer is free again
                    host = stage[:m].numpy()
                else:
                    host = out[a:a + m].numpy()
                got = f.readinto(memoryview(host))
                if got != m:
                    raise ShardError(f"{path}: section {s['name']!r} is truncated")
                if verify:
                    crc = zlib.crc32(host, crc)
                if to_gpu:
                    out[a:a + m].copy_(stage[:m], non_blocking=True)
            if verify and (crc & 0xFFFFFFFF) != s["crc32"]:
                raise ShardError(f"{path}: section {s['name']!r} failed its CRC-32")
            arrays[s["name"]] = out.view(dt).view(s["shape"]) if s["nbytes"] else torch.empty(s["shape"], dtype=dt, device=out.device)
    if on_gpu:
        torch.cuda.current_stream(dev).synchronize()
    return footer["meta"], arrays, blobs

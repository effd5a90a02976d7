"""Host-side document table of a device collection: row -> (chunk id, text, metadata), chunk id -> row.

The reference keeps these columns inside the SeekDB server (`collection.add(ids, embeddings, metadatas,
documents)`, seekdb_connector.py:78-82) and gets them back per hit (`:136-139`). Here they sit next to the
device-resident corpus, laid out so that the BASELINE configs (10 M chunks) are usable:

  * ids and texts are COLUMNS — one growable UTF-8 byte blob + an int64 end-offset array each — not a Python
    object per row; a bulk `add` of a million nodes joins and appends, it does not loop over rows;
  * chunk ids that the store generates itself (uuid4, seekdb_connector.py:75) are produced in bulk from one
    `os.urandom` call with numpy bit twiddling;
  * metadata keeps the caller's dict objects by reference (a list `extend`) for rows added in this process, and
    stays an unparsed JSON column for rows that came from a shard file until a hit asks for one (`get`);
  * the id -> row dictionary is only built when something needs it (delete, or a caller supplying its own ids);
  * the filter columns (tag word, chat group; context/context.py:74-111) are computed per batch by
    `RowTagger.encode_batch` and stored in the shard, so a load never re-interns 10 M rows.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np


class StringColumn:
    """Nullable UTF-8 strings: blob + end offsets (+ a null byte per row). Row i is blob[ends[i-1]:ends[i]]."""

    def __init__(self):
        self.blob = bytearray()
        self._ends = np.zeros(1024, dtype=np.int64)
        self._null = np.zeros(1024, dtype=np.uint8)
        self.n = 0

    def __len__(self):
        return self.n

    def _reserve(self, need: int):
        if need > self._ends.shape[0]:
            cap = max(need, 2 * self._ends.shape[0])
            self._ends = np.concatenate([self._ends[:self.n], np.zeros(cap - self.n, dtype=np.int64)])
            self._null = np.concatenate([self._null[:self.n], np.zeros(cap - self.n, dtype=np.uint8)])

    def extend(self, values: Sequence[Optional[str]]):
        m = len(values)
        if m == 0:
            return
        self._reserve(self.n + m)
        enc = [b"" if v is None else (v if isinstance(v, bytes) else str(v).encode("utf-8")) for v in values]
        lens = np.fromiter(map(len, enc), dtype=np.int64, count=m)
        base = len(self.blob)
        self.blob += b"".join(enc)
        np.cumsum(lens, out=self._ends[self.n:self.n + m])
        self._ends[self.n:self.n + m] += base
        self._null[self.n:self.n + m] = np.fromiter((v is None for v in values), dtype=np.uint8, count=m)
        self.n += m

    def extend_fixed_ascii(self, data: bytes, width: int):
        """m rows of exactly `width` ASCII bytes each, already concatenated (generated uuids)."""
        m = len(data) // width
        self._reserve(self.n + m)
        base = len(self.blob)
        self.blob += data
        self._ends[self.n:self.n + m] = base + width * np.arange(1, m + 1, dtype=np.int64)
        self._null[self.n:self.n + m] = 0
        self.n += m

    def truncate(self, n: int):
        if n < self.n:
            self.n = n
            del self.blob[int(self._ends[n - 1]) if n else 0:]

    def get_bytes(self, i: int) -> Optional[bytes]:
        if not 0 <= i < self.n:
            raise IndexError(i)
        if self._null[i]:
            return None
        lo = int(self._ends[i - 1]) if i else 0
        return bytes(self.blob[lo:int(self._ends[i])])

    def get(self, i: int) -> Optional[str]:
        b = self.get_bytes(i)
        return None if b is None else b.decode("utf-8")

    # -- persistence: three raw arrays ------------------------------------------------------
    def arrays(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        return (np.frombuffer(bytes(self.blob), dtype=np.uint8), self._ends[:self.n].copy(), self._null[:self.n].copy())

    @classmethod
    def from_arrays(cls, blob: np.ndarray, ends: np.ndarray, null: np.ndarray) -> "StringColumn":
        c = cls()
        c.blob = bytearray(np.ascontiguousarray(blob, dtype=np.uint8).tobytes())
        c.n = int(ends.shape[0])
        c._ends = np.ascontiguousarray(ends, dtype=np.int64).copy()
        c._null = np.ascontiguousarray(null, dtype=np.uint8).copy()
        if c.n and (int(c._ends[-1]) != len(c.blob) or (np.diff(c._ends) < 0).any() or c._ends[0] < 0):
            raise ValueError("string column offsets do not match the blob")
        if c._ends.shape[0] == 0:
            c._ends, c._null = np.zeros(1024, dtype=np.int64), np.zeros(1024, dtype=np.uint8)
        return c


class _ColumnView:
    """Read-only sequence over one column of the table (`table.ids[row]`, `len(table.texts)`, iteration)."""

    def __init__(self, table: "HostDocTable", getter):
        self._t, self._get = table, getter

    def __len__(self):
        return len(self._t)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._get(j) for j in range(*i.indices(len(self._t)))]
        if i < 0:
            i += len(self._t)
        return self._get(i)

    def __iter__(self):
        return (self._get(j) for j in range(len(self._t)))

    def __eq__(self, other):
        return list(self) == list(other)


_HEX = np.frombuffer(b"0123456789abcdef", dtype=np.uint8)
_LAZY = object()   # a metadata cell that still lives in the JSON column


def bulk_uuid4(m: int) -> bytes:
    """m canonical uuid4 strings (36 ASCII bytes each, concatenated): one urandom call + numpy formatting
    (the reference calls uuid.uuid4() per node, seekdb_connector.py:75)."""
    raw = np.frombuffer(os.urandom(16 * m), dtype=np.uint8).reshape(m, 16).copy()
    raw[:, 6] = (raw[:, 6] & 0x0F) | 0x40   # version 4
    raw[:, 8] = (raw[:, 8] & 0x3F) | 0x80   # RFC 4122 variant
    hexd = np.empty((m, 32), dtype=np.uint8)
    hexd[:, 0::2] = _HEX[raw >> 4]
    hexd[:, 1::2] = _HEX[raw & 15]
    out = np.full((m, 36), ord("-"), dtype=np.uint8)
    out[:, 0:8], out[:, 9:13], out[:, 14:18] = hexd[:, 0:8], hexd[:, 8:12], hexd[:, 12:16]
    out[:, 19:23], out[:, 24:36] = hexd[:, 16:20], hexd[:, 20:32]
    return out.tobytes()


class HostDocTable:
    """Row -> (chunk id, text, metadata); chunk id -> row for deletes. Column storage (module docstring)."""

    def __init__(self):
        self._ids = StringColumn()
        self._texts = StringColumn()
        self._metas: List[object] = []              # dict / None by reference, or _LAZY (see _meta_json)
        self._meta_json: Optional[StringColumn] = None
        self._row_of: Optional[Dict[str, int]] = None   # built on first use
        self.deleted: set = set()

    def __len__(self):
        return self._ids.n

    @property
    def live(self) -> int:
        return len(self) - len(self.deleted)

    # column views (tests and callers index them like lists)
    @property
    def ids(self):
        return _ColumnView(self, self._ids.get)

    @property
    def texts(self):
        return _ColumnView(self, self._texts.get)

    @property
    def metas(self):
        return _ColumnView(self, self._meta)

    @property
    def row_of(self) -> Dict[str, int]:
        if self._row_of is None:
            blob, ends = bytes(self._ids.blob), self._ids._ends
            d, lo = {}, 0
            dele = self.deleted
            for r in range(len(self)):
                hi = int(ends[r])
                if r not in dele:
                    d[blob[lo:hi].decode("utf-8")] = r
                lo = hi
            self._row_of = d
        return self._row_of

    def _meta(self, row: int):
        if not 0 <= row < len(self):
            raise IndexError(row)
        m = self._metas[row]
        if m is _LAZY:
            raw = self._meta_json.get_bytes(row)
            m = None if raw is None else json.loads(raw.decode("utf-8"))
            self._metas[row] = m
        return m

    def validate_new(self, ids: Optional[Sequence[str]], m: int, texts, metas):
        """Everything that can reject an append, checked BEFORE the device or the table is touched."""
        for what, col in (("ids", ids), ("texts", texts), ("metas", metas)):
            if col is not None and len(col) != m:
                raise ValueError(f"{len(col)} {what} for {m} rows")
        if ids is not None:
            if len(set(ids)) != m:
                raise ValueError("duplicate id inside the batch")
            known = self.row_of
            for i in ids:
                if i in known:
                    raise ValueError(f"duplicate id {i!r}")

    def append(self, ids: Optional[Sequence[str]], texts: Optional[Sequence[Optional[str]]],
               metas: Optional[Sequence[Optional[dict]]], m: Optional[int] = None) -> List[str]:
        """Append m rows; `ids=None` generates uuid4s in bulk. All-or-nothing: raises before any column changes.
        Returns the ids."""
        if m is None:
            m = len(ids if ids is not None else texts if texts is not None else metas)
        self.validate_new(ids, m, texts, metas)
        start = len(self)
        if ids is None:
            data = bulk_uuid4(m)
            self._ids.extend_fixed_ascii(data, 36)
            s = data.decode("ascii")
            ids = [s[36 * j:36 * j + 36] for j in range(m)]
            if self._row_of is not None:
                self._row_of.update(zip(ids, range(start, start + m)))
        else:
            ids = [str(i) for i in ids]
            self._ids.extend(ids)
            self._row_of.update(zip(ids, range(start, start + m)))   # validate_new built it
        self._texts.extend(texts if texts is not None else [None] * m)
        self._metas.extend(metas if metas is not None else [None] * m)
        return ids

    def delete(self, ids: Iterable[str]) -> List[int]:
        rows = []
        known = self.row_of
        for i in ids:
            r = known.pop(i, None)
            if r is not None and r not in self.deleted:
                self.deleted.add(r)
                rows.append(r)
        return rows

    def get(self, row: int) -> Tuple[str, Optional[str], Optional[dict]]:
        return self._ids.get(row), self._texts.get(row), self._meta(row)

    # -- persistence ------------------------------------------------------------------------
    def to_arrays(self) -> Dict[str, np.ndarray]:
        """Raw sections for shardio: id / text / metadata columns + the tombstoned rows."""
        out = {}
        for name, col in (("ids", self._ids), ("texts", self._texts)):
            b, e, z = col.arrays()
            out[f"doc_{name}_blob"], out[f"doc_{name}_ends"], out[f"doc_{name}_null"] = b, e, z
        mj = StringColumn()
        cache: Dict[int, Optional[bytes]] = {}
        cells = []
        for r, m in enumerate(self._metas):
            if m is _LAZY:
                cells.append(self._meta_json.get_bytes(r))
            elif m is None:
                cells.append(None)
            else:
                k = id(m)
                if k not in cache:
                    cache[k] = json.dumps(m, ensure_ascii=False, separators=(",", ":")).encode("utf-8")
                cells.append(cache[k])
        mj.extend(cells)
        b, e, z = mj.arrays()
        out["doc_metas_blob"], out["doc_metas_ends"], out["doc_metas_null"] = b, e, z
        out["doc_deleted"] = np.array(sorted(self.deleted), dtype=np.int64)
        return out

    @classmethod
    def from_arrays(cls, arrays: Dict[str, np.ndarray], n: int) -> "HostDocTable":
        t = cls()
        cols = {}
        for name in ("ids", "texts", "metas"):
            cols[name] = StringColumn.from_arrays(arrays[f"doc_{name}_blob"], arrays[f"doc_{name}_ends"],
                                                  arrays[f"doc_{name}_null"])
            if cols[name].n != n:
                raise ValueError(f"doc table column {name!r} has {cols[name].n} rows, expected {n}")
        t._ids, t._texts, t._meta_json = cols["ids"], cols["texts"], cols["metas"]
        t._metas = [_LAZY] * n
        t.deleted = set(int(r) for r in arrays["doc_deleted"].tolist())
        return t

    @classmethod
    def from_lists(cls, ids, texts, metas, deleted=()) -> "HostDocTable":
        """Shards written before the column format kept one JSON blob of three lists."""
        t = cls()
        t._ids.extend([str(i) for i in ids])
        t._texts.extend(list(texts))
        t._metas = list(metas)
        t.deleted = set(int(r) for r in deleted)
        return t

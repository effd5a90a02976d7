"""Full-text (BM25) collections: chunk texts -> term ids -> device postings (K2).

The reference names a full-text index per collection (``generate_fulltext_index_name``,
super_rag/utils/utils.py:10-11) and a ``fulltext_search`` recall
(super_rag/schema/view_models.py:1043-1047,1077) but ships neither index nor runner. This module is
the host side of that recall: a process-global registry of collections keyed by the same name the
vector store uses, tokenisation, and the mapping from result rows back to (text, metadata).

Tokenisation: the reference's tokenizer (``super_rag.utils.tokenizer.get_default_tokenizer``,
tiktoken cl100k_base, utils/tokenizer.py:7-9) when it is importable and its vocabulary file is
available; otherwise a deterministic word-hash tokenizer (lower-cased ``\\w+`` tokens, CRC32 mod
vocab). Either way a term is an int in [0, vocab) and scoring is the numeric contract of
include/srag_b200.h (oracle/bm25.py). Scoring and top-k run on the GPU only.
"""
from __future__ import annotations

import re
import threading
import zlib
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import numpy as np
import torch

from ._lib import SragError
from .bm25 import Bm25Index
from .models import DocumentWithScore

_WORD = re.compile(r"\w+", re.UNICODE)
HASH_VOCAB = 1 << 18
CL100K_VOCAB = 100_352


def hash_tokenizer(vocab: int = HASH_VOCAB) -> Callable[[str], List[int]]:
    def encode(text: str) -> List[int]:
        return [zlib.crc32(w.encode("utf-8")) % vocab for w in _WORD.findall(text.lower())]
    return encode


def default_tokenizer():
    """(encode, vocab): the reference tokenizer if usable here, else the word-hash tokenizer."""
    try:
        from super_rag.utils.tokenizer import get_default_tokenizer  # type: ignore
        enc = get_default_tokenizer()
        enc("probe")
        return enc, CL100K_VOCAB
    except Exception:
        return hash_tokenizer(), HASH_VOCAB


class FulltextCollection:
    """Chunk texts of one collection with a BM25 index in HBM. Rebuilt lazily after appends."""

    def __init__(self, name: str, tokenizer: Optional[Callable[[str], List[int]]] = None, vocab: Optional[int] = None,
                 device: Optional[str] = None, apply_filter: bool = False):
        if tokenizer is None:
            tokenizer, vocab = default_tokenizer()
        if vocab is None:
            raise ValueError("vocab is required with a custom tokenizer")
        self.name, self.tokenizer, self.vocab = name, tokenizer, int(vocab)
        self.device = torch.device(device) if device else None
        self.texts: List[Optional[str]] = []
        self.metas: List[Optional[dict]] = []
        self._tokens: List[np.ndarray] = []
        self._index: Optional[Bm25Index] = None
        # metadata pre-filter (same tag word / group encoding as the vector store); off unless asked for, like
        # the vector connector's ctx["apply_filter"]
        from .vectorstore import RowTagger
        self.apply_filter = bool(apply_filter)
        self.tagger = RowTagger()
        self._tags: List[int] = []
        self._groups: List[int] = []
        self._tags_dev: Optional[torch.Tensor] = None
        self._groups_dev: Optional[torch.Tensor] = None
        self.lock = threading.Lock()

    def __len__(self):
        return len(self.texts)

    def _dev(self) -> torch.device:
        if self.device is None:
            if not torch.cuda.is_available():
                raise SragError("fulltext_search needs a CUDA device: there is no CPU fallback")
            self.device = torch.device("cuda", torch.cuda.current_device())
        return self.device

    def add(self, texts: Sequence[str], metas: Optional[Sequence[Optional[dict]]] = None) -> range:
        metas = list(metas) if metas is not None else [None] * len(texts)
        with self.lock:
            start = len(self.texts)
            for t, m in zip(texts, metas):
                ids = np.asarray([i for i in self.tokenizer(t or "") if 0 <= i < self.vocab], dtype=np.int64)
                self.texts.append(t)
                self.metas.append(m)
                self._tokens.append(ids)
                tag, group = self.tagger.encode(m)
                self._tags.append(tag - (1 << 32) if tag & 0x80000000 else tag)
                self._groups.append(group)
            self._index = None
            return range(start, len(self.texts))

    def index(self) -> Bm25Index:
        with self.lock:
            if self._index is None:
                dev = self._dev()
                lens = np.fromiter((len(t) for t in self._tokens), dtype=np.int64, count=len(self._tokens))
                doc_ptr = np.zeros(len(lens) + 1, dtype=np.int64)
                np.cumsum(lens, out=doc_ptr[1:])
                flat = np.concatenate(self._tokens) if len(self._tokens) else np.zeros(0, dtype=np.int64)
                self._index = Bm25Index.from_tokens(torch.from_numpy(flat).to(dev), torch.from_numpy(doc_ptr).to(dev),
                                                    self.vocab)
                self._tags_dev = torch.tensor(self._tags, dtype=torch.int32).to(dev)
                self._groups_dev = torch.tensor(self._groups, dtype=torch.int32).to(dev)
            return self._index

    # -- persistence: token ids, texts and metadata (the tokenizer itself is code, passed again on load) --
    def save(self, path: str) -> None:
        from . import shardio
        with self.lock:
            lens = np.fromiter((len(t) for t in self._tokens), dtype=np.int64, count=len(self._tokens))
            doc_ptr = np.zeros(len(lens) + 1, dtype=np.int64)
            np.cumsum(lens, out=doc_ptr[1:])
            flat = np.concatenate(self._tokens) if len(self._tokens) else np.zeros(0, dtype=np.int64)
            shardio.write_shard(path, "fulltext", {"name": self.name, "vocab": self.vocab, "n_docs": len(self.texts)},
                                {"tokens": torch.from_numpy(flat), "doc_ptr": torch.from_numpy(doc_ptr)},
                                {"docs": {"texts": self.texts, "metas": self.metas}})

    @classmethod
    def load(cls, path: str, tokenizer: Optional[Callable[[str], List[int]]] = None, vocab: Optional[int] = None,
             name: Optional[str] = None, device: Optional[str] = None) -> "FulltextCollection":
        from . import shardio
        meta, a, blobs = shardio.read_shard(path, expect_kind="fulltext")
        if tokenizer is not None and vocab is None:
            vocab = int(meta["vocab"])
        col = cls(name or meta["name"], tokenizer, vocab, device)
        if col.vocab != int(meta["vocab"]):
            raise shardio.ShardError(f"{path}: vocabulary size {meta['vocab']} != tokenizer's {col.vocab}")
        tokens, doc_ptr = a["tokens"].numpy(), a["doc_ptr"].numpy()
        col.texts, col.metas = list(blobs["docs"]["texts"]), list(blobs["docs"]["metas"])
        if len(doc_ptr) != len(col.texts) + 1 or (len(doc_ptr) and int(doc_ptr[-1]) != len(tokens)):
            raise shardio.ShardError(f"{path}: token table does not match the doc table")
        col._tokens = [tokens[doc_ptr[i]:doc_ptr[i + 1]].copy() for i in range(len(col.texts))]
        for m in col.metas:
            tag, group = col.tagger.encode(m)
            col._tags.append(tag - (1 << 32) if tag & 0x80000000 else tag)
            col._groups.append(group)
        return col

    def query_terms(self, query: Optional[str], keywords: Optional[Iterable[str]]) -> List[int]:
        """Custom keywords replace the query text (FulltextSearchParams.keywords, view_models.py:1045-1047)."""
        text = " ".join(keywords) if keywords else (query or "")
        return list(self.tokenizer(text))

    def search(self, query: Optional[str], top_k: int, keywords: Optional[Iterable[str]] = None,
               filter: Optional[dict] = None) -> List[DocumentWithScore]:
        """`filter`: a document in the grammar of ContextManager._create_combined_filter (context/context.py:74-111);
        honoured when the collection was created with apply_filter=True, ignored otherwise."""
        if top_k <= 0 or not self.texts:
            return []
        terms = self.query_terms(query, keywords)
        index = self.index()
        kw = {}
        if filter and self.apply_filter:
            mask, group = self.tagger.compile(filter)
            dev = index.device
            if mask != 0xFFFFFFFF:
                kw["doc_tags"] = self._tags_dev
                kw["q_tag_mask"] = torch.tensor([mask - (1 << 32) if mask & 0x80000000 else mask], dtype=torch.int32).to(dev)
            if group >= 0:
                kw["doc_group"] = self._groups_dev
                kw["q_group"] = torch.tensor([group], dtype=torch.int32).to(dev)
        scores, ids = index.search([terms], min(int(top_k), 256), **kw)
        out: List[DocumentWithScore] = []
        for s, r in zip(scores[0].tolist(), ids[0].tolist()):
            if r < 0:
                break
            out.append(DocumentWithScore(text=self.texts[r], metadata=self.metas[r], score=s))
        return out


_REGISTRY: Dict[str, FulltextCollection] = {}
_REGISTRY_LOCK = threading.Lock()


def get_or_create_collection(name: str, **kwargs) -> FulltextCollection:
    with _REGISTRY_LOCK:
        if name not in _REGISTRY:
            _REGISTRY[name] = FulltextCollection(name, **kwargs)
        return _REGISTRY[name]


def get_collection(name: str) -> Optional[FulltextCollection]:
    return _REGISTRY.get(name)


def delete_collection(name: str) -> None:
    with _REGISTRY_LOCK:
        _REGISTRY.pop(name, None)

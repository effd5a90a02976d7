"""On-disk shard format: layout, CRC and truncation detection on the CPU; collection /
BM25 index round trips with bit-identical search results on the GPU."""
import json
import os

import numpy as np
import pytest
import torch

from super_rag_b200 import shardio


def _arrays():
    g = torch.Generator().manual_seed(3)
    return {
        "rows": torch.randn(37, 16, generator=g).to(torch.bfloat16),
        "row_scale": torch.tensor([1.0, float("nan"), 0.5] + [2.0] * 34),
        "ptr": torch.arange(9, dtype=torch.int64),
        "empty": torch.zeros(0, dtype=torch.int32),
    }


def test_round_trip_cpu(tmp_path):
    p = str(tmp_path / "a.shard")
    arrays = _arrays()
    blob = {"ids": ["a", "b"], "texts": ["héllo", None], "metas": [{"k": 1}, None]}
    footer = shardio.write_shard(p, "vector", {"n": 37, "dim": 16}, arrays, {"docs": blob})
    assert not os.path.exists(p + ".tmp")
    assert all(s["offset"] % shardio.ALIGN == 0 for s in footer["sections"])
    meta, got, blobs = shardio.read_shard(p, expect_kind="vector")
    assert meta == {"n": 37, "dim": 16} and blobs["docs"] == blob
    for k, t in arrays.items():
        assert got[k].dtype == t.dtype and tuple(got[k].shape) == tuple(t.shape)
        assert torch.equal(got[k].view(torch.uint8) if t.numel() else got[k], t.view(torch.uint8) if t.numel() else t), k
    assert torch.isnan(got["row_scale"][1])           # tombstones survive as bit patterns
    with pytest.raises(shardio.ShardError):
        shardio.read_shard(p, expect_kind="bm25")


def test_corruption_and_truncation_are_detected(tmp_path):
    p = str(tmp_path / "a.shard")
    footer = shardio.write_shard(p, "vector", {}, _arrays())
    raw = bytearray(open(p, "rb").read())
    sec = next(s for s in footer["sections"] if s["name"] == "rows")
    raw[sec["offset"] + 5] ^= 0x40
    bad = str(tmp_path / "bad.shard")
    open(bad, "wb").write(raw)
    with pytest.raises(shardio.ShardError, match="CRC"):
        shardio.read_shard(bad)
    shardio.read_shard(bad, verify=False)              # layout itself is intact
    cut = str(tmp_path / "cut.shard")
    open(cut, "wb").write(open(p, "rb").read()[:-11])
    with pytest.raises(shardio.ShardError):
        shardio.read_shard(cut)
    open(cut, "wb").write(b"not a shard at all, just bytes.....")
    with pytest.raises(shardio.ShardError, match="magic"):
        shardio.read_shard(cut)


def test_multi_chunk_streaming(tmp_path, monkeypatch):
    monkeypatch.setattr(shardio, "CHUNK_BYTES", 1000)   # force many staging chunks
    p = str(tmp_path / "a.shard")
    t = torch.arange(10_007, dtype=torch.int32)
    shardio.write_shard(p, "bm25", {"x": 1}, {"t": t})
    _, got, _ = shardio.read_shard(p)
    assert torch.equal(got["t"], t)
    assert json.loads(json.dumps(shardio.read_footer(p)))["kind"] == "bm25"


@pytest.mark.gpu
def test_collection_and_bm25_round_trip_gpu(tmp_path):
    from super_rag_b200 import fulltext
    from super_rag_b200.bm25 import Bm25Index
    from super_rag_b200.vectorstore import B200Client
    rng = np.random.default_rng(8)
    n, d = 3000, 128
    C = rng.standard_normal((n, d)).astype(np.float32)
    client = B200Client()
    client.delete_collection("t_shard")
    col = client.create_collection("t_shard", d)
    ids = col.append(torch.from_numpy(C), texts=[f"t{r}" for r in range(n)], metas=[{"row": r} for r in range(n)])
    col.delete([ids[5], ids[77]])
    Q = torch.from_numpy(C[[5, 77, 100, 2999]] + 0.01)
    s0, i0 = col.search_batch(Q, 7)
    p = str(tmp_path / "vec.shard")
    client.save_collection("t_shard", p)
    client.delete_collection("t_shard")
    col2 = client.load_collection(p)
    s1, i1 = col2.search_batch(Q, 7)
    assert torch.equal(i0, i1) and torch.equal(s0.view(torch.int32), s1.view(torch.int32))
    assert 5 not in i1[0].tolist() and col2.docs.live == n - 2 and col2.docs.get(100)[1] == "t100"
    assert col2.delete([ids[5]]) == 0 and col2.delete([ids[100]]) == 1            # tombstones and the id map survive
    more = col2.append(torch.from_numpy(C[:3]))                                  # appends keep working after a load
    assert len(more) == 3 and col2.n == n + 3
    client.delete_collection("t_shard")

    dev = torch.device("cuda:0")
    vocab = 2000
    dl = rng.integers(10, 60, 5000)
    pz = 1.0 / np.arange(1, vocab + 1)
    tokens = rng.choice(vocab, size=int(dl.sum()), p=pz / pz.sum()).astype(np.int64)
    doc_ptr = np.concatenate([[0], np.cumsum(dl)]).astype(np.int64)
    idx = Bm25Index.from_tokens(torch.from_numpy(tokens).to(dev), torch.from_numpy(doc_ptr).to(dev), vocab, id_base=40)
    queries = [rng.integers(0, vocab, 5).tolist() for _ in range(16)]
    bs0, bi0 = idx.search(queries, 10)
    pb = str(tmp_path / "bm25.shard")
    idx.save(pb)
    idx2 = Bm25Index.load(pb, dev)
    bs1, bi1 = idx2.search(queries, 10)
    assert torch.equal(bi0, bi1) and torch.equal(bs0.view(torch.int32), bs1.view(torch.int32))
    assert idx2.id_base == 40 and idx2.stats.avgdl == idx.stats.avgdl and (idx2.stats.df == idx.stats.df).all()
    assert torch.equal(idx2.post_off, idx.post_off) and torch.equal(idx2.slice_ptr, idx.slice_ptr)   # used as stored: no rebuild
    # a shard as round 1 wrote it (int32 documents, unpadded) still loads: it is packed again
    cd, ci = idx.canonical_postings()
    pb1 = str(tmp_path / "bm25_r1.shard")
    shardio.write_shard(pb1, "bm25", {"n_docs": idx.n_docs, "vocab": vocab, "id_base": 40, "stats_n_docs": idx.stats.n_docs,
                                      "avgdl": float(idx.stats.avgdl), "bank_ordered": False, "block_docs": 2048},
                        {"term_ptr": idx.term_ptr, "post_doc": cd, "post_imp": ci, "dl": idx.dl,
                         "df": torch.from_numpy(np.ascontiguousarray(idx.stats.df, dtype=np.int64))})
    idx3 = Bm25Index.load(pb1, dev)
    bs3, bi3 = idx3.search(queries, 10)
    assert torch.equal(bi0, bi3) and torch.equal(bs0.view(torch.int32), bs3.view(torch.int32))
    assert torch.equal(idx3.post_off, idx.post_off)

    enc = fulltext.hash_tokenizer(4096)
    ft = fulltext.FulltextCollection("t_ft_shard", tokenizer=enc, vocab=4096)
    ft.add([f"alpha beta w{r % 50} gamma{r % 7}" for r in range(500)], [{"row": r} for r in range(500)])
    want = [(d.text, d.score) for d in ft.search("w3 gamma2", 5)]
    pf = str(tmp_path / "ft.shard")
    ft.save(pf)
    ft2 = fulltext.FulltextCollection.load(pf, tokenizer=enc)
    assert [(d.text, d.score) for d in ft2.search("w3 gamma2", 5)] == want and want

"""CPU-only tests of the host layer above the C ABI: the drop-in connector / adaptor, the node pack,
sharding helpers and the "no CPU fallback, no oracle in the product" rules. No kernel is launched."""
import asyncio
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = "/root/reference"


def _run(coro):
    return asyncio.run(coro)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "super_rag_b200")
    pat = re.compile(r"^\s*(from\s+oracle\b|import\s+oracle\b|from\s+\.\.?oracle\b)", re.M)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f), encoding="utf-8").read()
                assert not pat.search(src), f"{f} imports the oracle"
                assert "/root/reference" not in src, f"{f} reads the reference at run time"


def test_no_cpu_fallback_is_loud():
    from super_rag_b200 import SragError, ops
    c = torch.zeros(8, 64, dtype=torch.bfloat16)
    s = torch.ones(8)
    with pytest.raises(SragError):
        ops.topk_search(c, s, c[:2], s[:2], 3)
    with pytest.raises(SragError):
        ops.ingest_rows(torch.zeros(4, 64))
    with pytest.raises(SragError):
        ops.fuse(torch.zeros(2, 1, 4, dtype=torch.int64), None, ops.FUSE_RRF, 2)
    if not torch.cuda.is_available():
        from super_rag_b200.vectorstore import B200VectorStoreConnector
        from super_rag_b200.models import TextNode
        con = B200VectorStoreConnector({"collection": "t_nofallback", "vector_size": 64})
        with pytest.raises(SragError):
            con.add([TextNode(text="x", embedding=[0.0] * 64)])
        con.delete_collection()


def test_adaptor_contract():
    from super_rag_b200.vectorstore import B200VectorStoreConnector, VectorStoreConnectorAdaptor
    a = VectorStoreConnectorAdaptor("b200", {"collection": "t_adaptor"})
    assert isinstance(a.connector, B200VectorStoreConnector)
    assert a.connector.store is a.connector and a.connector.vector_size == 1024 and a.connector.distance == "cosine"
    with pytest.raises(KeyError):                      # seekdb_connector.py:39 ctx["collection"]
        VectorStoreConnectorAdaptor("b200", {})
    with pytest.raises(ValueError):                    # seekdb_connector.py:95-96
        a.connector.delete()
    if not os.path.isdir(REFERENCE):
        with pytest.raises(ValueError):                # connector.py:14-15
            VectorStoreConnectorAdaptor("qdrant", {"collection": "x"})
    a.connector.delete_collection()


def test_doc_table_and_tombstone_bookkeeping():
    from super_rag_b200.doctable import HostDocTable
    t = HostDocTable()
    assert t.append(["a", "b", "c"], ["ta", "tb", "tc"], [None, {"k": 1}, None]) == ["a", "b", "c"]
    with pytest.raises(ValueError):
        t.append(["a"], ["dup"], [None])
    with pytest.raises(ValueError):                    # a duplicate inside the batch, and nothing was appended
        t.append(["x", "x"], ["1", "2"], [None, None])
    with pytest.raises(ValueError):                    # ragged columns (zip() used to truncate them silently)
        t.append(["y", "z"], ["only one"], [None, None])
    assert len(t) == 3 and "x" not in t.row_of and "y" not in t.row_of
    assert t.delete(["b", "zz"]) == [1] and t.delete(["b"]) == [] and t.live == 2
    assert t.get(2) == ("c", "tc", None) and t.get(1) == ("b", "tb", {"k": 1})
    assert t.ids[0] == "a" and list(t.texts) == ["ta", "tb", "tc"] and len(t.metas) == 3


def test_doc_table_columns_bulk_uuid_and_round_trip():
    """Column storage: generated ids are valid, distinct uuid4s; arrays -> table round trip keeps every cell, parses
    metadata lazily and keeps tombstones."""
    import uuid
    from super_rag_b200.doctable import HostDocTable, bulk_uuid4
    raw = bulk_uuid4(1000).decode("ascii")
    ids = [raw[36 * i:36 * i + 36] for i in range(1000)]
    assert len(set(ids)) == 1000
    for i in ids[:50]:
        u = uuid.UUID(i)
        assert u.version == 4 and u.variant == uuid.RFC_4122 and str(u) == i
    t = HostDocTable()
    shared = {"indexer": "pdf", "n": 1}
    got = t.append(None, ["héllo", None, "", "x" * 1000], [shared, shared, None, {"chat_id": "c1"}])
    assert len(got) == 4 and t.get(0) == (got[0], "héllo", shared) and t.get(1)[1] is None and t.get(2)[1] == ""
    t.append(["mine"], ["t"], None)
    assert t.delete([got[1]]) == [1]
    t2 = HostDocTable.from_arrays(t.to_arrays(), len(t))
    assert len(t2) == 5 and t2.deleted == {1}
    for r in range(5):
        assert t2.get(r) == t.get(r)
    assert t2.row_of["mine"] == 4 and got[1] not in t2.row_of
    assert t2.append(None, None, None, m=3) and len(t2) == 8 and t2.get(7)[1:] == (None, None)


def test_row_tagger_encode_batch_matches_per_row_encode():
    from super_rag_b200.vectorstore import RowTagger, MISSING_BIT
    metas = [None, {}, {"indexer": "a"}, {"indexer": "b", "chat_id": 7}, {"chat_id": "7"}, {"indexer": "a"}]
    metas += [metas[3]] * 3
    t1, t2 = RowTagger(), RowTagger()
    tags, groups = t1.encode_batch(metas)
    ref = [t2.encode(m) for m in metas]
    assert [int(x) & 0xFFFFFFFF for x in tags] == [r[0] for r in ref] and groups.tolist() == [r[1] for r in ref]
    assert t1.state() == t2.state() and (int(tags[0]) & 0xFFFFFFFF) == MISSING_BIT
    assert RowTagger.from_state(t1.state()).compile({"chat_id": 7}) == t1.compile({"chat_id": 7})


def test_merge_union_matches_reference_golden(golden):
    from super_rag_b200 import nodeflow_pack as n
    from super_rag_b200.models import DocumentWithScore as D
    n.register()
    entry = n.NODE_RUNNER_REGISTRY["merge"]
    runner = entry["runner"]
    assert entry["input_model"] is n.MergeInput or not n.USING_REFERENCE_NODEFLOW
    for c in golden["merge_union"]:
        i = c["inputs"]
        mk = lambda k: [D(**d) for d in i[k]]
        ui = n.MergeInput(merge_strategy="union", deduplicate=c["deduplicate"], vector_search_docs=mk("v"),
                          fulltext_search_docs=mk("f"), graph_search_docs=mk("g"), summary_search_docs=mk("s"),
                          vision_search_docs=mk("vi"))
        out, so = _run(runner.run(ui, None))
        assert [d.model_dump() for d in out.docs] == c["docs"] and so == c["system_output"]
    with pytest.raises(n.ValidationError, match="Unknown merge strategy: bogus"):
        _run(runner.run(n.MergeInput(merge_strategy="bogus"), None))
    # defaults of the reference MergeInput are preserved (merge.py:13-14)
    assert n.MergeInput().merge_strategy == "union" and n.MergeInput().deduplicate is True


def test_node_pack_registers_reference_shaped_nodes():
    from super_rag_b200 import nodeflow_pack as n
    n.register()
    for t in ("fulltext_search", "merge", "vector_search"):
        e = n.NODE_RUNNER_REGISTRY[t]
        assert set(e) == {"runner", "input_model", "output_model"}
        assert "docs" in e["output_model"].model_json_schema()["properties"]      # {{ nodes.X.output.docs }}
    f = n.FulltextSearchInput()
    assert f.top_k == 5 and f.keywords is None and f.collection_ids == []
    # unknown collection / no collection ids degrade to [] like the vector node
    si = n.SystemInput(query="q", user="u")
    out, _ = _run(n.NODE_RUNNER_REGISTRY["fulltext_search"]["runner"].run(n.FulltextSearchInput(), si))
    assert out.docs == []
    out, _ = _run(n.NODE_RUNNER_REGISTRY["fulltext_search"]["runner"].run(
        n.FulltextSearchInput(collection_ids=["nope"]), si))
    assert out.docs == []


def test_hash_tokenizer_and_collection_bookkeeping():
    from super_rag_b200 import fulltext
    enc = fulltext.hash_tokenizer(1000)
    assert enc("Hello, hello WORLD") == [enc("hello")[0]] * 2 + enc("world") and all(0 <= t < 1000 for t in enc("a b c"))
    col = fulltext.FulltextCollection("t_ft", tokenizer=enc, vocab=1000)
    assert list(col.add(["alpha beta", "beta gamma gamma"], [None, {"m": 1}])) == [0, 1]
    assert len(col) == 2 and col._tokens[1].tolist() == enc("beta gamma gamma")
    assert col.query_terms("alpha", ["beta", "gamma"]) == enc("beta gamma")     # keywords replace the query
    assert col.search("alpha", 0) == []


def test_shard_range_and_exchange_keys():
    import numpy as np
    from super_rag_b200.sharded import pack_keys_torch, shard_range, unpack_keys_torch
    for n, g in [(10_000_000, 8), (7, 3), (0, 2), (5, 8)]:
        edges = [shard_range(n, g, r) for r in range(g)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(edges[r][1] == edges[r + 1][0] for r in range(g - 1))
    # exchange keys: orderable(score) << 32 | (2^32-1 - id); round trip, empty slots, unsigned order == (score desc, id asc)
    s = torch.tensor([[1.5, -0.0, float("-inf"), 0.0, -2.25, 1.5]], dtype=torch.float32)
    i = torch.tensor([[7, 3, -1, 9, (1 << 32) - 3, 6]], dtype=torch.int64)
    keys = pack_keys_torch(s, i)
    s2, i2 = unpack_keys_torch(keys)
    assert torch.equal(i2, i) and torch.equal(s2[0, [0, 3, 4, 5]], s[0, [0, 3, 4, 5]])
    assert float(s2[0, 1]) == 0.0 and float(s2[0, 2]) == float("-inf") and int(keys[0, 2]) == 0
    ku = [int(x) for x in keys.numpy().view(np.uint64)[0]]
    order = sorted(range(6), key=lambda j: -ku[j])
    assert order == [5, 0, 1, 3, 4, 2]                   # 1.5 (id 6, then 7); the two zeros by id (-0.0 == +0.0); -2.25; empty


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference checkout not present on this machine")
def test_install_connector_into_unmodified_reference():
    """VECTOR_DB_TYPE=b200 resolves inside the reference's own ContextManager (context/context.py:8-12)."""
    sys.path.insert(0, REFERENCE)
    try:
        import super_rag.vectorstore.connector as ref_conn
        from super_rag.context.context import ContextManager
        from super_rag_b200 import nodeflow_pack as n, vectorstore
        original = ref_conn.VectorStoreConnectorAdaptor
        try:
            assert n.install_connector() and n.install_connector()           # idempotent
            cm = ContextManager("t_ctx", None, "b200", {"collection": "t_ctx", "vector_size": 64})
            assert isinstance(cm.adaptor.connector, vectorstore.B200VectorStoreConnector)
            with pytest.raises(ValueError):                                   # other types reach the reference adaptor
                ContextManager("t_ctx", None, "no-such-store", {"collection": "t_ctx"})
            # the reference's own filter builder now also runs for type b200 (it returns None for non-seekdb stores)
            flt = cm._create_combined_filter(["vector"], "chat-9")
            assert flt == {"and": [{"or": [{"indexer": {"$in": ["vector"]}}, {"indexer": {"$exists": False}}]},
                                   {"chat_id": "chat-9"}]}
            assert cm.vectordb_type == "b200" and cm._create_combined_filter(None, None) is None
            tagger = vectorstore.RowTagger()
            tagger.encode({"indexer": "vector", "chat_id": "chat-9"})
            assert tagger.compile(flt) == (vectorstore.MISSING_BIT | 1, 1)
        finally:
            ContextManager._create_combined_filter = ContextManager._create_combined_filter._original
            ref_conn.VectorStoreConnectorAdaptor = original
            sys.modules["super_rag.context.context"].VectorStoreConnectorAdaptor = original
            vectorstore._REFERENCE_ADAPTOR = None
    finally:
        sys.path.remove(REFERENCE)


def test_unmodified_reference_connector_over_the_pyseekdb_stub():
    """SURVEY §8(c) last row: the reference's own SeekDBVectorStoreConnector.search + ContextManager.query, imported
    unmodified from /root/reference with only `pyseekdb` stubbed (tests/ref_stub.py: exact-cosine collection.query),
    return the oracle's ranked rows wrapped in the reference's DocumentWithScore. Skipped where the reference is absent."""
    import numpy as np
    import oracle
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
    import ref_stub
    if not ref_stub.reference_available():
        pytest.skip("needs /root/reference")
    seek, ContextManager, models = ref_stub.load_reference_connector()
    rng = np.random.default_rng(3)
    n, d, k = 500, 32, 5
    C = rng.standard_normal((n, d)).astype(np.float32)
    C[400] = C[7]
    C[33] = 0
    name = "t_ref_stub"
    ref_stub.STATE.collections.pop(name, None)
    cm = ContextManager(name, embedding_model=None, vectordb_type="seekdb", vectordb_ctx={"collection": name, "vector_size": d})
    con = cm.adaptor.connector
    assert type(con).__name__ == "SeekDBVectorStoreConnector" and type(con).__module__.startswith("super_rag.")
    con.create_collection(vector_size=d)
    from super_rag_b200.models import TextNode
    ids = con.add([TextNode(text=f"t{r}", metadata={"row": r}, embedding=C[r].tolist()) for r in range(n)])
    assert len(ids) == n
    for qi in (7, 33, 123):
        docs = cm.query("hello", score_threshold=0.2, topk=k, vector=C[qi].tolist(), index_types=["vector"], chat_id="c")
        want_s, want_i = oracle.exact_topk(C, C[qi:qi + 1], k, "cosine")
        assert [x.metadata["row"] for x in docs] == want_i[0].tolist()
        assert np.allclose([x.score for x in docs], want_s[0], atol=1e-6)
        assert isinstance(docs[0], models.DocumentWithScore) and docs[0].text == f"t{want_i[0, 0]}"
    assert [x.metadata["row"] for x in cm.query("q", topk=2, vector=C[7].tolist())] == [7, 400]    # duplicates: lower row first
    con.delete(ids=[ids[7]])
    assert [x.metadata["row"] for x in cm.query("q", topk=1, vector=C[7].tolist())] == [400]
    with pytest.raises(ValueError):
        con.delete()
    con.delete_collection()

"""GPU tests of the write path behind the store faces: incremental device append, upsert,
tombstone deletes (both metrics), compaction, persistence, the Cassandra row loader, and readers
racing a writer.  Every answer is checked against the oracle on the rows that are live."""
import threading

import numpy as np
import pytest

import oracle
from githubrepostorag_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def _expect(q, emb16, live_ids, k, mode="default", keep=None):
    s, r = oracle.query_vectorised(q, emb16, k, mode=mode, keep=keep)
    return [live_ids[int(i)] for i in r], s


def _ids_of(res):
    return res.ids, np.asarray(res.similarities, np.float32)


def _query(st, q, k, flt=None):
    from githubrepostorag_b200.types import MetadataFilter, MetadataFilters, VectorStoreQuery
    f = MetadataFilters(filters=[MetadataFilter(key=a, value=b) for a, b in flt.items()]) if flt else None
    return st.query(VectorStoreQuery(query_embedding=q.tolist(), similarity_top_k=k, filters=f))


@pytest.mark.parametrize("metric", ["cosine", "ip"])
def test_incremental_add_upsert_delete_compact(metric):
    from githubrepostorag_b200.vector_store import B200VectorStore
    mode = oracle.MODE_COSINE if metric == "cosine" else oracle.MODE_DOT
    rng = np.random.default_rng(5)
    d, k = 64, 7
    emb = rng.standard_normal((6000, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    ids = [S.row_id(i) for i in range(len(emb))]
    metas = [{"repo": f"r{i % 3}", "namespace": "default"} for i in range(len(emb))]
    st = B200VectorStore(dim=d, metric=metric)
    q = rng.standard_normal(d, dtype=np.float32)

    state = {}          # id -> (vector, meta): the live rows, in insertion order of their latest version

    def apply_add(lo, hi):
        st.add_embeddings(ids[lo:hi], emb[lo:hi], [f"t{i}" for i in range(lo, hi)], metas[lo:hi])
        for i in range(lo, hi):
            state.pop(ids[i], None)
            state[ids[i]] = (emb[i], metas[i])

    def check(flt=None):
        live = list(state)
        e16 = np.stack([state[i][0] for i in live]).astype(np.float16)
        keep = None if flt is None else np.array([all(state[i][1].get(a) == b for a, b in flt.items()) for i in live])
        want_ids, want_s = _expect(q, e16, live, k, mode, keep)
        got_ids, got_s = _ids_of(_query(st, q, k, flt))
        assert got_ids == want_ids, (flt, got_ids, want_ids)
        scale = 1.0 if metric == "cosine" else float(np.abs(want_s).max() + 1)
        assert np.allclose(got_s, want_s, rtol=1e-5, atol=2e-7 * scale)
        assert len(st) == len(live)

    apply_add(0, 500); check()
    cap0 = st._shard.capacity
    apply_add(500, 1700); check(); check({"repo": "r1"})           # crosses the first capacity: buffers grow in place
    assert st._shard.capacity > cap0 and st._shard.n_rows == 1700
    # upsert: new vectors for existing ids (old rows are tombstoned on the device)
    emb[100:140] = rng.standard_normal((40, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    apply_add(100, 140); check()
    assert st._shard.n_rows == 1740
    # delete the current best hits: they must never come back
    top_ids, _ = _ids_of(_query(st, q, k))
    for rid in top_ids[:4]:
        st.delete(rid); state.pop(rid)
    check(); check({"repo": "r0", "namespace": "default"})
    # mass delete triggers compaction (tombstones outnumber live rows)
    for i in range(200, 1700):
        if ids[i] in state:
            st.delete(ids[i]); state.pop(ids[i])
    check()
    assert st._rows.n_dead == 0 and st._shard.n_rows == len(state)
    apply_add(1700, 6000); check(); check({"repo": "r2"})
    st.close()


def test_save_load_roundtrip_and_cassandra_loader(tmp_path):
    from githubrepostorag_b200.migrate import load_cassandra_rows
    from githubrepostorag_b200.vector_store import B200VectorStore
    rng = np.random.default_rng(8)
    n, d = 3000, 384
    vec = rng.standard_normal((n, d), dtype=np.float32)
    cql_rows = [{"row_id": f"id-{i}", "attributes_blob": None, "body_blob": f"body {i}", "vector": vec[i].tolist(),
                 "metadata_s": {"namespace": "default", "repo": f"repo{i % 7}", "scope": "chunk", "extra": "x"}}
                for i in range(n)]
    cql_rows.append({"row_id": "broken", "body_blob": "no vector", "vector": None, "metadata_s": {}})
    st = B200VectorStore(dim=d)
    assert load_cassandra_rows(st, cql_rows, batch=700) == n + 1 and len(st) == n
    q = vec[123] + 0.01 * rng.standard_normal(d, dtype=np.float32)
    a = _query(st, q, 6, {"repo": "repo4"})
    assert a.ids[0] == "id-123" and a.nodes[0].text == "body 123" and a.nodes[0].metadata["extra"] == "x"
    st.delete("id-123")
    path = tmp_path / "embeddings.npz"
    st.save(str(path))
    st2 = B200VectorStore.load(str(path))
    b1, b2 = _query(st, q, 6, {"repo": "repo4"}), _query(st2, q, 6, {"repo": "repo4"})
    assert b1.ids == b2.ids and "id-123" not in b2.ids and b1.similarities == b2.similarities
    assert len(st2) == n - 1 and st2.get_by_document_id("id-7").page_content == "body 7"
    st2.add_embeddings(["id-new"], q[None, :], ["fresh"], [{"repo": "repo4"}])
    assert _query(st2, q, 1, {"repo": "repo4"}).ids == ["id-new"]
    # the streaming (directory) format: rows and codes go through a memory map block by block
    B200VectorStore.SAVE_BLOCK_ROWS = 1000                         # several blocks even for this small table
    st.save(str(tmp_path / "table_dir"))
    assert sorted(p.name for p in (tmp_path / "table_dir").iterdir()) == ["codes.npy", "meta.json", "rows.jsonl", "rows16.npy"]
    st3 = B200VectorStore.load(str(tmp_path / "table_dir"))
    b3 = _query(st3, q, 6, {"repo": "repo4"})
    assert b3.ids == b1.ids and b3.similarities == b1.similarities and b3.nodes[0].text == b1.nodes[0].text
    assert len(st3) == n - 1 and st3.get_by_document_id("id-7").metadata["repo"] == "repo0"
    st.close(); st2.close(); st3.close()


def test_readers_race_a_writer():
    from githubrepostorag_b200.vector_store import B200VectorStore
    rng = np.random.default_rng(3)
    d = 32
    emb = rng.standard_normal((20_000, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    st = B200VectorStore(dim=d)
    st.add_embeddings([S.row_id(i) for i in range(2000)], emb[:2000])
    errors, stop = [], threading.Event()

    def reader(seed):
        r = np.random.default_rng(seed)
        try:
            while not stop.is_set():
                j = int(r.integers(0, 2000))
                res = _query(st, emb[j], 1)
                assert res.ids == [S.row_id(j)] and abs(res.similarities[0] - 1.0) < 1e-6
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    ts = [threading.Thread(target=reader, args=(s,)) for s in range(6)]
    [t.start() for t in ts]
    for lo in range(2000, 20_000, 1500):                     # writer: appends force re-binds and growth
        st.add_embeddings([S.row_id(i) for i in range(lo, lo + 1500)], emb[lo:lo + 1500])
        assert _query(st, emb[lo + 7], 1).ids == [S.row_id(lo + 7)]
    stop.set()
    [t.join() for t in ts]
    assert not errors, errors[:2]
    st.close()


def test_list_valued_membership_and_in_filters():
    """`topics` / `imports` / `labels` are lists; a filter on them means membership.
    Also the IN operator on a coded key, and membership edges in adjacent()."""
    from githubrepostorag_b200.types import FilterOperator, MetadataFilter, MetadataFilters, VectorStoreQuery
    from githubrepostorag_b200.vector_store import B200VectorStore
    rng = np.random.default_rng(11)
    n, d, k = 5000, 48, 9
    emb = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    vocab = ["kafka", "jms", "redis", "grpc", "auth", "sql"]
    metas = []
    for i in range(n):
        t = [vocab[j] for j in range(len(vocab)) if rng.random() < 0.2]
        md = {"namespace": "default", "repo": f"r{i % 6}", "imports": [f"lib{i % 40}", "os"]}
        if t:
            md["topics"] = t
        metas.append(md)
    ids = [S.row_id(i) for i in range(n)]
    st = B200VectorStore(dim=d)
    st.add_embeddings(ids[:3000], emb[:3000], None, metas[:3000])
    q = rng.standard_normal(d, dtype=np.float32)

    def expect(pred):
        keep = np.array([pred(m) for m in metas[:len(st)]])
        s, r = oracle.query_vectorised(q, emb[:len(st)].astype(np.float16), k, keep=keep)
        return [ids[int(i)] for i in r]

    assert _query(st, q, k, {"topics": "kafka"}).ids == expect(lambda m: "kafka" in m.get("topics", ()))
    assert _query(st, q, k, {"topics": "kafka", "repo": "r2"}).ids == \
        expect(lambda m: "kafka" in m.get("topics", ()) and m["repo"] == "r2")
    assert _query(st, q, k, {"topics": "grpc", "imports": "lib7"}).ids == \
        expect(lambda m: "grpc" in m.get("topics", ()) and "lib7" in m["imports"])
    assert _query(st, q, k, {"topics": "no-such-topic"}).ids == []
    st.add_embeddings(ids[3000:], emb[3000:], None, metas[3000:])          # masks are rebuilt after a write
    assert _query(st, q, k, {"topics": "kafka"}).ids == expect(lambda m: "kafka" in m.get("topics", ()))
    f = MetadataFilters(filters=[MetadataFilter(key="repo", value=["r1", "r4", "zzz"], operator=FilterOperator.IN),
                                 MetadataFilter(key="topics", value="sql", operator=FilterOperator.CONTAINS)])
    got = st.query(VectorStoreQuery(query_embedding=q.tolist(), similarity_top_k=k, filters=f))
    assert got.ids == expect(lambda m: m["repo"] in ("r1", "r4") and "sql" in m.get("topics", ()))
    assert got.nodes[0].metadata["topics"] == metas[int(got.ids[0][4:])]["topics"]      # lists come back as lists
    adj = st.adjacent({("topics", "auth"), ("repo", "r5")}, q.tolist(), k=k, filter={"namespace": "default"})
    assert [c.id for c in adj] == expect(lambda m: "auth" in m.get("topics", ()) or m["repo"] == "r5")
    st.delete(got.ids[0])
    assert got.ids[0] not in _query(st, q, k, {"topics": "sql"}).ids
    st.close()

"""GPU tests of the filter vocabulary behind the store faces: the ingest writer's per-scope key set (`owner`,
`symbol`, ... -- ingest/src/app/services/vector_write_service.py:28-34,58-64), OR / nested conditions, IN on
list-valued keys, stores over a row-sharded range, and concurrent masked searches.  Every answer is checked against
the oracle's `prefilter_mask` + `query_vectorised` on the same rows."""
import threading

import numpy as np
import pytest

import oracle
from githubrepostorag_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def _corpus(seed, n, d):
    rng = np.random.default_rng(seed)
    emb = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    vocab = ["kafka", "jms", "redis", "grpc", "auth", "sql"]
    metas = []
    for i in range(n):
        md = {"namespace": "default" if i % 5 else "acme", "repo": f"r{i % 7}", "owner": f"o{i % 3}",
              "module": f"m{i % 23}", "file_path": f"src/f{i % 211}.py", "symbol": f"fn{i % 97}",
              "language": ["python", "java", "go"][i % 3], "scope": "chunk", "branch": "main" if i % 10 else "dev",
              "component_kind": "service" if i % 4 else "standalone"}
        t = [vocab[j] for j in range(len(vocab)) if rng.random() < 0.25]
        if t:
            md["topics"] = t
        md["imports"] = [f"lib{i % 31}", "os"]
        if i % 2:
            md["labels"] = [f"L{i % 4}"]
        metas.append(md)
    return rng, emb, metas, [S.row_id(i) for i in range(n)]


def _expect(q, emb, metas, ids, k, flt, condition="and"):
    keep = oracle.prefilter_mask(metas, flt, condition=condition)
    s, r = oracle.query_vectorised(q, emb.astype(np.float16), k, keep=keep)
    return [ids[int(i)] for i in r], s


@pytest.mark.parametrize("scope,table", [("repo", "embeddings_repo"), ("chunk", "embeddings"), ("file", "embeddings_file")])
def test_every_writer_key_filters_on_its_table(scope, table):
    """`filter={"owner": ...}` / `{"symbol": ...}` (and every other key the writer indexes or always keeps) return
    the oracle's rows on the table of that scope -- through the LangChain face, the LlamaIndex face and adjacent()."""
    from githubrepostorag_b200.migrate import store_for_table
    from githubrepostorag_b200.types import VectorStoreQuery
    from githubrepostorag_b200.vector_store import filter_keys_for_scope
    n, d, k = 4000, 64, 8
    rng, emb, metas, ids = _corpus(17, n, d)
    st = store_for_table(table, dim=d)
    assert st.scope == scope
    st.add_embeddings(ids, emb, [f"t{i}" for i in range(n)], metas)
    q = rng.standard_normal(d, dtype=np.float32)
    coded, multi = filter_keys_for_scope(scope)
    for key in coded:
        flt = {key: metas[11][key]}
        want_ids, want_s = _expect(q, emb, metas, ids, k, flt)
        got = st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=flt)
        assert [i for _, _, i in got] == want_ids, key
        assert np.allclose([2 * s - 1 for _, s, _ in got], want_s, rtol=1e-5, atol=2e-6), key
    for key in multi:
        flt = {key: next(m for m in metas if key in m)[key][0]}
        assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=flt)] == \
            _expect(q, emb, metas, ids, k, flt)[0], key
    flt = {"owner": "o1", "symbol": "fn5", "namespace": "default"}
    res = st.query(VectorStoreQuery(query_embedding=q.tolist(), similarity_top_k=k, filters=flt))
    assert res.ids == _expect(q, emb, metas, ids, k, flt)[0]
    # edges on the new keys (GraphRetrieverFactory.py:107-145 lists owner / symbol / language as edges)
    adj = st.adjacent({("owner", "o2"), ("symbol", "fn3")}, q.tolist(), k=k, filter={"namespace": "default"})
    want = _expect(q, emb, metas, ids, k, {"namespace": "default", "__any_of__": [[{"owner": "o2"}, {"symbol": "fn3"}]]})[0]
    assert [c.id for c in adj] == want
    assert st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter={"owner": "nobody"}) == []
    st.close()


def test_or_nested_and_in_filters_match_the_oracle():
    from githubrepostorag_b200.types import (FilterCondition, FilterOperator, MetadataFilter, MetadataFilters,
                                             VectorStoreQuery)
    from githubrepostorag_b200.vector_store import B200VectorStore
    n, d, k = 6000, 48, 10
    rng, emb, metas, ids = _corpus(23, n, d)
    st = B200VectorStore(dim=d)
    st.add_embeddings(ids, emb, None, metas)
    q = rng.standard_normal(d, dtype=np.float32)
    F = MetadataFilter

    def ask(filters):
        return st.query(VectorStoreQuery(query_embedding=q.tolist(), similarity_top_k=k, filters=filters)).ids

    # flat OR of equalities on coded and list-valued keys
    got = ask(MetadataFilters([F("repo", "r3"), F("owner", "o1"), F("topics", "sql", FilterOperator.CONTAINS)],
                              condition=FilterCondition.OR))
    assert got == _expect(q, emb, metas, ids, k, {"repo": "r3", "owner": "o1", "topics": "sql"}, condition="or")[0]
    # AND of an equality and a nested OR
    got = ask(MetadataFilters([F("namespace", "default"),
                               MetadataFilters([F("symbol", "fn7"), F("language", ["go", "rust"], FilterOperator.IN)],
                                               condition=FilterCondition.OR)]))
    want = _expect(q, emb, metas, ids, k, {"namespace": "default",
                                           "__any_of__": [[{"symbol": "fn7"}, {"language": ("in", ["go", "rust"])}]]})[0]
    assert got == want
    # IN on a list-valued key: the row's list shares a member with the IN list
    got = ask(MetadataFilters([F("topics", ["kafka", "grpc", "nope"], FilterOperator.IN), F("branch", "main")]))
    assert got == _expect(q, emb, metas, ids, k, {"topics": ("in", ["kafka", "grpc", "nope"]), "branch": "main"})[0]
    assert ask(MetadataFilters([F("topics", ["nope"], FilterOperator.IN)])) == []
    # OR whose alternatives are all unknown matches nothing; an unknown alternative next to a known one is ignored
    assert ask(MetadataFilters([F("repo", "zz"), F("nokey", "1")], condition=FilterCondition.OR)) == []
    got = ask(MetadataFilters([F("repo", "zz"), F("repo", "r1")], condition=FilterCondition.OR))
    assert got == _expect(q, emb, metas, ids, k, {"repo": "r1"})[0]
    # the same OR through the LangChain face (dict form) and after a write (mask cache invalidated)
    flt = {"__any_of__": [[{"owner": "o0"}, {"labels": "L3"}]]}
    assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=flt)] == \
        _expect(q, emb, metas, ids, k, flt)[0]
    st.delete(ids[int(_expect(q, emb, metas, ids, 1, flt)[0][0][4:])])
    metas2 = list(metas)
    metas2[int(_expect(q, emb, metas, ids, 1, flt)[0][0][4:])] = {}
    assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=flt)] == \
        _expect(q, emb, metas2, ids, k, flt)[0]
    st.close()


def test_store_over_a_row_shard_with_offset():
    """A store whose device shard starts at global row `lo` (row-sharding): searches return global rows, while
    ids / text / metadata / stored vectors are looked up by LOCAL row."""
    from githubrepostorag_b200.index import MetadataDictionary
    from githubrepostorag_b200.vector_store import B200VectorStore
    n, d, k, lo = 5000, 64, 6, 123_456
    rng = np.random.default_rng(31)
    emb16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    repo = (np.arange(n) % 9).astype(np.int32)
    dic = MetadataDictionary(["repo"])
    dic.register_vocab("repo", [f"r{i}" for i in range(9)])
    st = B200VectorStore.from_arrays(emb16, repo[None, :].copy(), dic, id_fn=lambda r: S.row_id(lo + r),
                                     meta_fn=lambda r: {"repo": f"r{r % 9}"}, text_fn=lambda r: f"text {lo + r}",
                                     row_offset=lo)
    q = rng.standard_normal(d, dtype=np.float32)
    for flt, keep in ((None, None), ({"repo": "r4"}, repo == 4)):
        s_ref, r_ref = oracle.query_vectorised(q, emb16, k, keep=keep)
        got = st.similarity_search_with_embedding_id_by_vector(q.tolist(), k=k, filter=flt)
        assert [i for _, _, i in got] == [S.row_id(lo + int(r)) for r in r_ref]
        for (doc, vec, rid), r in zip(got, r_ref):
            assert np.array_equal(np.asarray(vec, np.float32), emb16[int(r)].astype(np.float32))
            assert doc.page_content == f"text {lo + int(r)}" and doc.metadata["repo"] == f"r{int(r) % 9}"
        assert np.allclose([d_.metadata["score"] for d_, _, _ in got], s_ref, rtol=1e-5, atol=2e-7)
    adj = st.adjacent({("repo", "r1"), ("repo", "r2")}, q.tolist(), k=k)
    s_ref, r_ref = oracle.query_vectorised(q, emb16, k, keep=(repo == 1) | (repo == 2))
    assert [c.id for c in adj] == [S.row_id(lo + int(r)) for r in r_ref]
    assert st.get([S.row_id(lo + 5)])[0].content == f"text {lo + 5}"
    st.close()


def test_concurrent_masked_searches_do_not_share_scratch():
    """The worker runs up to 10 jobs on a thread pool against ONE store (rag_worker/src/worker/worker.py:136-139,185),
    and the agent routinely filters on `topics` (agent_graph.py:230-233) -- a list-valued key, i.e. the device path
    with a row mask.  Threads searching under different masks at the same time must each get the serial answer."""
    from githubrepostorag_b200.vector_store import B200VectorStore
    n, d, k = 60_000, 96, 10
    rng, emb, metas, ids = _corpus(41, n, d)
    st = B200VectorStore(dim=d)
    st.add_embeddings(ids, emb, None, metas)
    topics = ["kafka", "jms", "redis", "grpc", "auth", "sql"]
    qs = rng.standard_normal((12, d), dtype=np.float32)
    jobs = [(qs[i], {"topics": topics[i % 6], **({"repo": f"r{i % 7}"} if i % 2 else {})}) for i in range(12)]
    serial = [[i for _, _, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=f)] for q, f in jobs]
    for j, (q, f) in enumerate(jobs[:4]):
        assert serial[j] == _expect(q, emb, metas, ids, k, f)[0]
    errors = []
    start = threading.Barrier(len(jobs))

    def run(j):
        q, f = jobs[j]
        try:
            start.wait()
            for _ in range(25):
                got = [i for _, _, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=f)]
                if got != serial[j]:
                    errors.append((j, got, serial[j]))
                    return
        except Exception as e:  # noqa: BLE001
            errors.append((j, repr(e)))

    ts = [threading.Thread(target=run, args=(j,)) for j in range(len(jobs))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors[:2]
    st.close()


def test_narrow_code_columns_equal_the_int32_matrix_and_widen_on_append():
    """Code columns are stored as narrow as their dictionaries allow (int8 / int16 / int32, the filter kernel is bound
    by their bytes).  Same survivors as the int32 matrix -- odd row counts (scalar tail), many filters in one pass,
    missing codes, caller row masks, the pass-bit masks the tensor-core plan reads -- and a column is widened in place
    when an append brings a code its type cannot hold."""
    import torch
    from githubrepostorag_b200.index import DeviceShard
    rng = np.random.default_rng(77)
    n, d, k = 70_003, 64, 12
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    codes = np.stack([rng.integers(-1, 4, n), rng.integers(-1, 2000, n), rng.integers(0, 2, n),
                      rng.integers(0, 70_000, n)]).astype(np.int32)          # int8, int16, int8, int32 columns
    wide = DeviceShard.from_host(rows16, codes, device=0, narrow=False)
    slim = DeviceShard.from_host(rows16, codes, device=0)
    assert wide.code_widths == [4, 4, 4, 4] and slim.code_widths == [1, 2, 1, 4]
    wants = [[0, -1, -1, -1], [-1, 7, -1, -1], [3, -1, 1, -1], [-1, -1, -1, 123], [2, 1999, 0, -1], [-1, -1, 1, -1],
             [1, -1, -1, -1], [-2, -1, -1, -1], [-1, 0, 0, -1], [-1, -1, 0, -1], [-1, 5, 1, -1], [0, 0, 0, 0]]
    extra = rng.random(n) < 0.4
    for batch_wants, masked in ((wants, False), (wants[:3], True), ([wants[5]] * 40, False)):
        b = len(batch_wants)
        filt = np.asarray(batch_wants, np.int32)
        q = torch.from_numpy(rng.standard_normal((b, d), dtype=np.float32)).cuda()
        res = []
        for sh in (wide, slim):
            masks = [sh.pack_row_mask(torch.from_numpy(extra).cuda())] * b if masked else None
            for algo in ((0, 2) if b == 40 else (0,)):
                s_, r_, v_ = sh.search_device(q, k, filt=filt, row_masks=masks, algo=algo)
                torch.cuda.synchronize()
                res.append((s_.clone(), r_.clone(), v_.clone()))
        for other in res[1:]:
            assert torch.equal(res[0][1], other[1]) and torch.equal(res[0][2], other[2])
            assert torch.equal(res[0][0].view(torch.int32), other[0].view(torch.int32))
        qn = q.cpu().numpy()
        for i in (0, b - 1):
            keep = np.ones(n, bool)
            for m, w in enumerate(batch_wants[i]):
                if w == -2:
                    keep[:] = False
                elif w >= 0:
                    keep &= codes[m] == w
            if masked:
                keep &= extra
            s_ref, r_ref = oracle.query_vectorised(qn[i], rows16, k, keep=keep)
            assert res[0][1][i, :len(r_ref)].cpu().numpy().tolist() == r_ref.tolist()
    # an append whose codes outgrow int8 / int16 widens those columns; old and new rows stay filterable
    more = rng.standard_normal((500, d), dtype=np.float32).astype(np.float16)
    more_codes = np.stack([np.full(500, 300), np.full(500, 40_000), np.zeros(500, int), np.arange(500)]).astype(np.int32)
    slim.append(more, more_codes)
    assert slim.code_widths == [2, 4, 1, 4] and slim.n_rows == n + 500
    allrows = np.concatenate([rows16, more])
    allcodes = np.concatenate([codes, more_codes], axis=1)
    qv = rng.standard_normal((1, d), dtype=np.float32)
    for want in ([300, -1, -1, -1], [-1, 40_000, 0, -1], [2, -1, 1, -1], [-1, 1999, -1, -1]):
        keep = np.ones(n + 500, bool)
        for m, w in enumerate(want):
            if w >= 0:
                keep &= allcodes[m] == w
        s_ref, r_ref = oracle.query_vectorised(qv[0], allrows, k, keep=keep)
        s_, r_, v_ = slim.search_host(qv, k, filt=np.asarray([want], np.int32))
        assert int(v_[0]) == len(r_ref) and r_[0, :len(r_ref)].tolist() == r_ref.tolist(), want
    wide.close()
    slim.close()


def test_a_code_every_row_carries_is_dropped_from_the_filter(monkeypatch):
    """The worker pins `namespace` on every request (agent_graph.py:198-239); in a single-namespace table that
    predicate is true of every row.  The plan drops it (histograms of the code columns, exact): same result bit for
    bit as with the elimination switched off, as the unfiltered search and as the oracle -- with the unfiltered plan's launches, no
    filter pass; deleted rows stay excluded; an append that ends the uniformity ends the elimination."""
    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.index import DeviceShard
    rng = np.random.default_rng(5)
    n, d, k = 60_000, 384, 10
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    codes = np.stack([np.zeros(n, np.int32), rng.integers(0, 40, n).astype(np.int32)])
    sh = DeviceShard.from_host(rows16, codes, device=0, capacity=n + 16)
    q = rng.standard_normal((1, d), dtype=np.float32)

    def run(filt):
        before = N.launch_count()
        s_, r_, v_ = sh.search_host(q, k, filt=None if filt is None else np.asarray([filt], np.int32))
        return s_.copy(), r_.copy(), v_.copy(), N.launch_count() - before

    s0, r0, v0, l0 = run(None)
    s1, r1, v1, l1 = run([0, -1])
    assert l1 == l0 and np.array_equal(r0, r1) and np.array_equal(s0.view(np.int32), s1.view(np.int32))
    s2, r2, v2, l2 = run([0, 7])                       # the second key still filters; one key left for the filter pass
    keep = codes[1] == 7
    s_ref, r_ref = oracle.query_vectorised(q[0], rows16, k, keep=keep)
    assert r2[0].tolist() == r_ref.tolist()
    monkeypatch.setenv("CODERAG_B200_DROP_TRIVIAL", "0")
    monkeypatch.setenv("CODERAG_B200_ZONE_MAPS", "0")          # (a code's row range covering the shard says the same thing)
    sh.reload_tuning()
    s3, r3, v3, l3 = run([0, -1])
    s4, r4, v4, _ = run([0, 7])
    assert l3 > l1 and np.array_equal(r3, r1) and np.array_equal(s3.view(np.int32), s1.view(np.int32))
    assert np.array_equal(r4, r2) and np.array_equal(s4.view(np.int32), s2.view(np.int32))
    monkeypatch.delenv("CODERAG_B200_DROP_TRIVIAL")
    monkeypatch.delenv("CODERAG_B200_ZONE_MAPS")
    sh.reload_tuning()
    # a deleted row never comes back, filter dropped or not
    top = int(r1[0, 0])
    sh.tombstone([top])
    _, r5, _, l5 = run([0, -1])
    live = np.ones(n, bool)
    live[top] = False
    assert l5 == l0 and r5[0].tolist() == oracle.query_vectorised(q[0], rows16, k, keep=live)[1].tolist()
    # one row of another namespace: the predicate constrains again, and that row is filtered out
    extra = (q[0] * 3).astype(np.float16)[None, :]     # would be the best hit
    sh.append(extra, np.asarray([[1], [7]], np.int32))
    _, r6, _, l6 = run([0, -1])          # (namespace 0 is still one run, rows [0, n): a streaming pass over it)
    assert l6 == l0 and n not in r6[0].tolist() and r6[0].tolist() == r5[0].tolist()
    monkeypatch.setenv("CODERAG_B200_ZONE_MAPS", "0")          # without row ranges the filter pass is back
    sh.reload_tuning()
    _, r6b, _, l6b = run([0, -1])
    monkeypatch.delenv("CODERAG_B200_ZONE_MAPS")
    sh.reload_tuning()
    assert l6b > l0 and r6b[0].tolist() == r6[0].tolist()
    _, r7, _, _ = run([1, -1])
    assert r7[0, 0] == n
    sh.close()


@pytest.mark.parametrize("d", [384, 768])
def test_one_search_over_the_union_of_alternative_filters(d):
    """crb_search_host_any_of: one query, a row passes when ANY alternative passes -- the adjacent() step of the graph
    traversal.  Equal to the oracle under the OR of the alternatives' masks, and (ids, scores, vectors) to the
    per-edge batch merged on the host; nested and overlapping sets, an alternative no row matches, one every row
    matches, the 32-alternative limit, deleted rows, both metrics."""
    from githubrepostorag_b200.index import DeviceShard
    rng = np.random.default_rng(d)
    n, k = 80_003, 8
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    repo = rng.integers(0, 6, n)
    module = repo * 50 + rng.integers(0, 50, n)                 # a module lies in one repo
    fpath = module * 200 + rng.integers(0, 200, n)              # a file in one module (60k values: an int32 column)
    codes = np.stack([np.zeros(n, np.int64), repo, module, fpath]).astype(np.int32)
    sh = DeviceShard.from_host(rows16, codes, device=0)
    assert sh.code_widths == [1, 1, 2, 4]
    q = rng.standard_normal(d, dtype=np.float32)
    r0 = int(np.argmax(rows16.astype(np.float32) @ q))           # a strong hit: its repo / module / file are the edges

    def mask_of(alt):
        keep = np.ones(n, bool)
        for m, w in enumerate(alt):
            if w == -2:
                keep[:] = False
            elif w >= 0:
                keep &= codes[m] == w
        return keep

    nested = [[0, repo[r0], -1, -1], [0, -1, module[r0], -1], [0, -1, -1, fpath[r0]]]
    cases = {
        "nested": nested,
        "overlapping": [[-1, 1, -1, -1], [-1, -1, module[r0], -1], [-1, 2, -1, -1], [-1, 1, int(module[r0]), -1]],
        "with an impossible one": nested + [[-2, -1, -1, -1]],
        "all impossible": [[-2, -1, -1, -1], [-1, -2, -1, -1]],
        "one matches everything": nested + [[0, -1, -1, -1]],
        "single": [[-1, 3, -1, -1]],
        "duplicates": [nested[1], nested[1], nested[2]],
        "32 files": [[-1, -1, -1, int(v)] for v in np.unique(fpath)[:32]],
    }
    for metric, mode in (("cosine", oracle.MODE_COSINE), ("ip", oracle.MODE_DOT)):
        for name, alts in cases.items():
            f = np.asarray(alts, np.int32)
            s_, r_, nv, vec = sh.search_host_any_of(q, f, k, metric=metric, vectors=True)
            keep = np.zeros(n, bool)
            for alt in alts:
                keep |= mask_of(alt)
            s_ref, r_ref = oracle.query_vectorised(q, rows16, k, keep=keep, mode=mode)
            assert nv == len(r_ref) and r_[:nv].tolist() == r_ref.tolist(), (name, metric)
            tol = 2e-6 * (1.0 if metric == "cosine" or not nv else float(np.abs(s_ref).max()) + 1.0)
            assert np.allclose(s_[:nv], s_ref, rtol=1e-5, atol=tol)
            assert np.array_equal(vec[:nv], rows16[r_[:nv]].astype(np.float32)) and not vec[nv:].any()
            # the per-edge batch, merged as the adapter merges: same rows, same score bits
            bs, br, bv = sh.search_host(np.repeat(q[None, :], len(alts), 0), k, metric=metric, filt=f)
            best = {}
            for i in range(len(alts)):
                for j in range(int(bv[i])):
                    best[int(br[i, j])] = bs[i, j]
            top = sorted(best.items(), key=lambda t: (-t[1], -t[0]))[:k]
            assert [r for r, _ in top] == r_[:nv].tolist(), (name, metric)
            assert np.array_equal(np.asarray([s for _, s in top], np.float32).view(np.int32), s_[:nv].view(np.int32)), (name, metric)
    with pytest.raises(RuntimeError, match="at most 32"):
        sh.search_host_any_of(q, np.asarray([[-1, -1, -1, int(v)] for v in np.unique(fpath)[:33]], np.int32), k)
    # a deleted row leaves the union
    _, r_, nv, _ = sh.search_host_any_of(q, np.asarray(nested, np.int32), k, vectors=True)
    sh.tombstone([int(r_[0])])
    _, r2, nv2, _ = sh.search_host_any_of(q, np.asarray(nested, np.int32), k, vectors=True)
    keep = mask_of(nested[0]) | mask_of(nested[1]) | mask_of(nested[2])
    keep[int(r_[0])] = False
    assert r2[:nv2].tolist() == oracle.query_vectorised(q, rows16, k, keep=keep)[1].tolist()
    sh.close()


def test_adjacent_takes_the_union_search_and_equals_the_per_edge_batch(monkeypatch):
    """B200VectorStore.adjacent(): coded-equality edges go out as one union search; edges on list-valued keys keep the
    per-edge batch.  Same Content list either way."""
    from githubrepostorag_b200.vector_store import B200VectorStore
    n, d, k = 6000, 64, 7
    rng, emb, metas, ids = _corpus(23, n, d)
    st = B200VectorStore(dim=d)
    st.add_embeddings(ids, emb, [f"t{i}" for i in range(n)], metas)
    q = rng.standard_normal(d, dtype=np.float32)
    calls = []
    real = B200VectorStore._search_any_of

    def spy(self, *a, **kw):
        out = real(self, *a, **kw)
        calls.append(out is not None)
        return out
    monkeypatch.setattr(B200VectorStore, "_search_any_of", spy)
    for edges, base in (([("repo", "r2"), ("module", "m5"), ("file_path", "src/f7.py")], {"namespace": "default"}),
                        ([("owner", "o1"), ("symbol", "fn3"), ("repo", "r6")], None),
                        ([("repo", "r2"), ("topics", "kafka")], {"namespace": "default"})):
        calls.clear()
        got = st.adjacent(edges, q.tolist(), k=k, filter=base)
        took_union = calls == [True]
        assert took_union == all(fld != "topics" for fld, _ in edges)
        monkeypatch.setattr(B200VectorStore, "MAX_ALTERNATIVES", 0)
        want = st.adjacent(edges, q.tolist(), k=k, filter=base)
        monkeypatch.setattr(B200VectorStore, "MAX_ALTERNATIVES", 32)
        assert [c.id for c in got] == [c.id for c in want] and [c.embedding for c in got] == [c.embedding for c in want]
        flt = {"__any_of__": [[{fld: val} for fld, val in edges]], **(base or {})}
        assert [c.id for c in got] == _expect(q, emb, metas, ids, k, flt)[0]
    st.close()


def _clustered(rng, n, d):
    """Rows ordered repository by repository (a module lies in one repository, a file in one module), plus a key that
    is scattered (`kind`) and one with a single value (`namespace`)."""
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    repo = rng.integers(0, 12, n)
    module = repo * 40 + rng.integers(0, 40, n)
    order = np.lexsort((module, repo))
    repo, module = repo[order], module[order]
    kind = rng.integers(0, 3, n)
    codes = np.stack([np.zeros(n, np.int64), repo, module, kind]).astype(np.int32)
    return rows16, codes


@pytest.mark.parametrize("d", [384, 1024])
def test_filters_on_contiguous_codes_are_range_scans(d, monkeypatch):
    """Per-code row ranges (taken at bind time): a filter whose codes each occupy one contiguous run is an UNFILTERED
    pass over the intersection of the runs -- no filter pass (launch counts), same bits as with the ranges switched off,
    same rows as the oracle; a filter that also names a scattered key is filtered inside the range; unions of runs that
    join up are one range; an append at the end widens a range and everything stays right."""
    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.index import DeviceShard
    rng = np.random.default_rng(d + 1)
    n, k = 120_007, 10
    rows16, codes = _clustered(rng, n, d)
    sh = DeviceShard.from_host(rows16, codes, device=0, capacity=n + 64)
    q = rng.standard_normal((2, d), dtype=np.float32)

    def keep_of(want, cds=codes):
        keep = np.ones(cds.shape[1], bool)
        for m, w in enumerate(want):
            if w >= 0:
                keep &= cds[m] == w
        return keep

    def run(fn):
        out = {}
        for z in ("1", "0"):
            monkeypatch.setenv("CODERAG_B200_ZONE_MAPS", z)
            sh.reload_tuning()
            before = N.launch_count()
            res = fn()
            out[z] = (res, N.launch_count() - before)
        monkeypatch.delenv("CODERAG_B200_ZONE_MAPS")
        sh.reload_tuning()
        return out

    plain = run(lambda: sh.search_host(q[:1], k))["1"][1]            # launches of an unfiltered host search
    r5, m5 = 5, int(codes[2][codes[1] == 5][100])
    cases = [([0, r5, -1, -1], True), ([-1, r5, m5, -1], True), ([-1, -1, m5, -1], True), ([-1, r5, -1, 2], False),
             ([-1, -1, m5, 1], False), ([-1, 3, m5, -1], True),      # a module of another repository: empty intersection
             ([-1, 11, -1, -1], True), ([-1, 0, -1, 0], False)]
    for want, exact in cases:
        for b in (1, 2):
            f = np.tile(np.asarray([want], np.int32), (b, 1))
            out = run(lambda: sh.search_host(q[:b], k, filt=f))
            (s1, r1, v1), l1 = out["1"]
            (s0, r0, v0), l0 = out["0"]
            assert np.array_equal(r1, r0) and np.array_equal(v1, v0) and np.array_equal(s1.view(np.int32), s0.view(np.int32)), want
            for i in range(b):
                s_ref, r_ref = oracle.query_vectorised(q[i], rows16, k, keep=keep_of(want))
                assert int(v1[i]) == len(r_ref) and r1[i, :len(r_ref)].tolist() == r_ref.tolist(), (want, b)
            if exact:
                assert l1 == plain and l0 >= plain, (want, b, l1, l0, plain)      # (the fused kernel is one launch too)
    # device path with a caller's row mask: filtered inside the range, never "exact"
    import torch
    extra = rng.random(n) < 0.5
    mask = sh.pack_row_mask(torch.from_numpy(extra).cuda())
    for want in ([-1, r5, -1, -1], [-1, r5, -1, 2]):
        f = np.asarray([want], np.int32)
        out = run(lambda: [t.cpu().numpy() for t in sh.search_device(torch.from_numpy(q[:1]).cuda(), k, filt=f, row_masks=[mask])])
        assert all(np.array_equal(a, b_) for a, b_ in zip(out["1"][0], out["0"][0]))
        s_ref, r_ref = oracle.query_vectorised(q[0], rows16, k, keep=keep_of(want) & extra)
        assert out["1"][0][1][0, :len(r_ref)].tolist() == r_ref.tolist()
    # unions: nested runs and neighbouring runs are one range (no filter pass); runs with a gap are filtered inside the hull
    m_next = int(codes[2][np.searchsorted(codes[2], m5, side="right")])        # the module stored right after m5
    for alts, exact in (([[-1, r5, -1, -1], [-1, -1, m5, -1]], True), ([[-1, -1, m5, -1], [-1, -1, m_next, -1]], True),
                        ([[-1, 2, -1, -1], [-1, 7, -1, -1]], False), ([[-1, 2, -1, -1], [-1, 3, -1, 1]], False),
                        ([[-1, 4, 9999, -1], [-1, 3, m5, -1]], True)):
        f = np.asarray(alts, np.int32)
        out = run(lambda: sh.search_host_any_of(q[0], f, k, vectors=True))
        (s1, r1, v1, e1), l1 = out["1"]
        (s0, r0, v0, e0), l0 = out["0"]
        assert v1 == v0 and np.array_equal(r1, r0) and np.array_equal(s1.view(np.int32), s0.view(np.int32)) and np.array_equal(e1, e0)
        keep = np.zeros(n, bool)
        for a in alts:
            keep |= keep_of(a)
        assert r1[:v1].tolist() == oracle.query_vectorised(q[0], rows16, k, keep=keep)[1].tolist(), alts
        if exact:
            assert l1 < l0, (alts, l1, l0)
    # an append lands at the end: repository 5's rows are no longer one run; results stay the oracle's
    more = rng.standard_normal((3, d), dtype=np.float32).astype(np.float16)
    more_codes = np.asarray([[0, 0, 0], [r5, r5, 0], [m5, m5, 1], [1, 1, 1]], np.int32)
    sh.append(more, more_codes)
    allrows, allcodes = np.concatenate([rows16, more]), np.concatenate([codes, more_codes], axis=1)
    for want in ([-1, r5, -1, -1], [-1, -1, m5, -1], [-1, 0, -1, -1], [-1, r5, -1, 1]):
        out = run(lambda: sh.search_host(q[:1], k, filt=np.asarray([want], np.int32)))
        (s1, r1, v1), _ = out["1"]
        assert np.array_equal(r1, out["0"][0][1])
        assert r1[0, :int(v1[0])].tolist() == oracle.query_vectorised(q[0], allrows, k, keep=keep_of(want, allcodes))[1].tolist()
    sh.close()


def test_store_clustered_by_its_keys_returns_the_same_documents():
    """B200VectorStore.compact(cluster_by=...): rows are re-ordered (a repository's rows, inside them a module's, become
    one run), ids / text / metadata follow, every face answers as before."""
    from githubrepostorag_b200.vector_store import B200VectorStore
    n, d, k = 5000, 64, 9
    rng, emb, metas, ids = _corpus(31, n, d)
    st = B200VectorStore(dim=d)
    st.add_embeddings(ids, emb, [f"t{i}" for i in range(n)], metas)
    st.delete(ids[7])
    q = rng.standard_normal(d, dtype=np.float32)
    flts = [None, {"repo": "r3"}, {"repo": "r3", "module": "m5"}, {"namespace": "default", "repo": "r2", "owner": "o1"},
            {"topics": "kafka", "repo": "r1"}, {"module": "m11"}]
    before = [[(i, round(s, 6)) for _, s, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=f)] for f in flts]
    adj_before = [c.id for c in st.adjacent([("repo", "r3"), ("module", "m5")], q.tolist(), k=k, filter={"namespace": "default"})]
    st.compact(cluster_by=("repo", "module"))
    sh = st._shard
    repo_col = sh.codes[st.dictionary.keys.index("repo")].cpu().numpy()
    assert (np.diff(repo_col) >= 0).all() and len(st) == n - 1
    after = [[(i, round(s, 6)) for _, s, i in st.similarity_search_with_score_id_by_vector(q.tolist(), k=k, filter=f)] for f in flts]
    assert after == before
    assert [c.id for c in st.adjacent([("repo", "r3"), ("module", "m5")], q.tolist(), k=k, filter={"namespace": "default"})] == adj_before
    live = np.ones(n, bool)
    live[7] = False
    assert [i for i, _ in after[1]] == _expect(q, emb[live], [m for j, m in enumerate(metas) if j != 7],
                                               [x for j, x in enumerate(ids) if j != 7], k, {"repo": "r3"})[0]
    doc = st.get_by_document_id(ids[11])
    assert doc.page_content == "t11" and doc.metadata["repo"] == metas[11]["repo"] and st.get([ids[7]]) == []
    st.add_embeddings(["late"], emb[:1], ["x"], [{"repo": "r3", "module": "m5"}])      # lands at the end: ranges widen
    got = st.similarity_search_with_score_id_by_vector(emb[0].tolist(), k=2, filter={"repo": "r3", "module": "m5"})
    assert "late" in [i for _, _, i in got]
    st.close()

"""CPU tests: the C-ABI library loads and exports what the header declares, and the host
logic above it (dictionary encoding, filter translation, store bookkeeping) behaves like the
reference's boundary.  No compute calls: there is no GPU here."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def lib():
    from githubrepostorag_b200 import _native, build
    build.build()
    return _native.load()


def _declared_symbols():
    text = (ROOT / "include" / "coderag_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(crb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound(lib):
    from githubrepostorag_b200 import _native
    declared = _declared_symbols()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/coderag_b200.h but not exported"
        assert name in _native.PROTOTYPES, f"{name} has no ctypes prototype"
    assert set(_native.PROTOTYPES) == set(declared)
    assert lib.crb_abi_version() == _native.ABI_VERSION


def test_argument_errors_are_reported_not_crashed(lib):
    from githubrepostorag_b200 import _native
    h = C.c_void_p()
    rc = lib.crb_index_create(C.byref(h), None, None, None, 10, 12, 0, 0, 0)   # dim % 8 != 0
    assert rc != 0 and b"dim" in lib.crb_last_error()
    rc = lib.crb_index_create(C.byref(h), None, None, None, 10, 16, 99, 0, 0)  # too many keys
    assert rc != 0 and b"n_keys" in lib.crb_last_error()
    rc = lib.crb_index_create(C.byref(h), None, None, None, 10, 16, 0, 2**32, 0)
    assert rc != 0
    with pytest.raises(RuntimeError, match="coderag_b200"):
        _native.check(rc)
    assert lib.crb_search(None, None, 1, 5, 0, None, None, None, None, None, None, 0, 0, None) != 0
    assert b"NULL" in lib.crb_last_error()
    assert lib.crb_merge_keys(None, 0, 1, 5, 0, None, 8, None, None, None, None, 0, None) != 0
    assert lib.crb_index_destroy(None) == 0
    assert lib.crb_workspace_bytes(None, 1, 1, 0) == 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from githubrepostorag_b200 import _native
    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "_LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _native.load()


def test_metadata_dictionary_codes_and_filters(lib):
    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.index import MetadataDictionary
    d = MetadataDictionary(["namespace", "repo", "branch"])
    codes = d.encode_rows([{"namespace": "default", "repo": "a"},
                           {"namespace": "acme", "repo": "a", "branch": "dev"},
                           {"repo": 7}])
    assert codes.dtype == np.int32 and codes.shape == (3, 3)
    assert codes[:, 0].tolist() == [0, 0, N.CODE_MISSING]
    assert codes[:, 1].tolist() == [1, 0, 0]
    assert codes[:, 2].tolist() == [N.CODE_MISSING, 1, N.CODE_MISSING]   # repo=7 stored as "7"
    assert d.encode_filter(None).tolist() == [N.FILT_ANY] * 3
    assert d.encode_filter({"repo": "a"}).tolist() == [N.FILT_ANY, 0, N.FILT_ANY]
    assert d.encode_filter({"repo": "7", "namespace": "acme"}).tolist() == [1, 1, N.FILT_ANY]
    assert d.encode_filter({"repo": "zzz"}).tolist() == [N.FILT_ANY, N.FILT_IMPOSSIBLE, N.FILT_ANY]
    assert d.encode_filter({"owner": "x"}).tolist() == [N.FILT_IMPOSSIBLE] * 3
    with pytest.raises(ValueError):
        MetadataDictionary([str(i) for i in range(N.MAX_KEYS + 1)])


def test_store_host_bookkeeping(lib):
    from githubrepostorag_b200.types import MetadataFilter, MetadataFilters, TextNode, VectorStoreQuery
    from githubrepostorag_b200.vector_store import B200VectorStore
    st = B200VectorStore(dim=8)
    assert st.stores_text and st.is_embedding_query and st.client is st
    nodes = [TextNode(text=f"t{i}", metadata={"repo": "a" if i % 2 else "b", "skip": None}, id_=f"n{i}",
                      embedding=[float(i)] * 8) for i in range(5)]
    assert st.add(nodes) == [f"n{i}" for i in range(5)]
    assert len(st) == 5
    st.add([TextNode(text="new", metadata={"repo": "c"}, id_="n2", embedding=[9.0] * 8)])   # upsert
    assert len(st) == 5 and st.get_by_document_id("n2").page_content == "new"
    st.delete("n0")
    assert len(st) == 4 and st.get_by_document_id("n0") is None
    st.delete("does-not-exist")
    with pytest.raises(ValueError):
        st.add_embeddings(["x"], np.zeros((1, 16), np.float32))
    f = MetadataFilters(filters=[MetadataFilter(key="repo", value="a"), MetadataFilter(key="namespace", value="n")])
    assert st._filters_to_dict(f) == {"repo": "a", "namespace": "n"}
    with pytest.raises(NotImplementedError):
        st.query(VectorStoreQuery(query_embedding=[0.0] * 8, similarity_top_k=1, node_ids=["n1"]))
    with pytest.raises(ValueError):
        st.query(VectorStoreQuery(query_embedding=None, similarity_top_k=1))


def test_synthetic_generator_is_block_stable():
    from githubrepostorag_b200 import synthetic as S
    full = S.corpus(1001, 5000, 16)
    part = S.corpus(1001, 5000, 16, lo=1234, hi=4321)
    assert full.dtype == np.float16 and np.array_equal(full[1234:4321], part)
    c_full = S.metadata_codes(1001, 5000)
    c_part = S.metadata_codes(1001, 5000, lo=100, hi=900)
    for k in S.FILTER_KEYS:
        assert np.array_equal(c_full[k][100:900], c_part[k])
    flt, sel = S.pick_filter(c_full, ("namespace", "component_kind"), 0.10)
    assert set(flt) == {"namespace", "component_kind"} and 0.0 < sel < 0.5
    md = S.metadata_dicts({k: v[:3] for k, v in c_full.items()})
    assert all(isinstance(v, str) for row in md for v in row.values())


def test_k1_d768_hot_loop_keeps_its_loads_together(lib):
    """Build-time guard (no GPU needed): the d=768 streaming kernel must keep its 12 row loads in one tight
    group.  When register pressure crosses the 128-register line that two resident CTAs allow, ptxas
    re-materialises row pointers between the loads and the kernel loses 7 % of HBM bandwidth
    (seen once; profiles/_ab/k1_ab.py measures it on a GPU)."""
    import shutil
    import subprocess
    from githubrepostorag_b200 import _native
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", str(_native.lib_path())], capture_output=True, text=True).stdout
    take, lines = False, []
    for ln in sass.splitlines():
        if "Function :" in ln:
            take = "k1_stream_kernelILi1ELi3ELi4ELi1ELi2ELb1E" in ln      # the exact-width (dim = 768) instantiation the plan dispatches
        elif take and "/*" in ln and ";" in ln:
            lines.append(ln)
    loads = [i for i, ln in enumerate(lines) if "LDG.E.NA.128" in ln]      # the streaming (no-allocate) row loads
    assert len(loads) >= 12, "d=768 kernel not found or has no 128-bit loads"
    first12 = loads[:12]
    assert first12[-1] - first12[0] <= 64, f"row loads are spread over {first12[-1] - first12[0]} instructions"


def test_tensor_core_kernels_are_tcgen05_tma_tmem_without_spills(lib):
    """Build-time evidence (no GPU needed): the K2 kernels issue tcgen05 MMAs fed by TMA and read their
    accumulators from TMEM -- not mma.sync / wmma -- and neither they nor the d=768 streaming kernel spill."""
    import re
    import shutil
    import subprocess
    from githubrepostorag_b200 import _native
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", str(_native.lib_path())], capture_output=True, text=True).stdout
    per_fn, cur = {}, None
    for ln in sass.splitlines():
        if "Function :" in ln:
            cur = ln.split("Function :")[1].strip()
            per_fn[cur] = []
        elif cur is not None:
            per_fn[cur].append(ln)
    # every TMA load is a plain (non-multicast) tile load; the only cluster multicast is tcgen05.commit (UTCBAR)
    assert not re.search(r"UTMALDG[.\w]*MULTICAST", sass), "a TMA multicast load is in the library"
    for tag in ("k2_gemm_topk_kernel", "k2_gemm_topk_pair_kernel"):
        names = [n for n in per_fn if tag in n]
        assert names, f"{tag} not in the library"
        body = "\n".join(per_fn[names[0]])
        assert "UTCHMMA" in body, f"{tag}: no tcgen05.mma (UTCHMMA)"
        assert "UTMALDG" in body, f"{tag}: no TMA tensor load (UTMALDG)"
        assert "LDTM" in body, f"{tag}: no TMEM load (LDTM)"
        assert not re.search(r"\bHMMA\b", body), f"{tag}: legacy mma.sync found"
    res = subprocess.run(["cuobjdump", "-res-usage", str(_native.lib_path())], capture_output=True, text=True).stdout
    fn = None
    for ln in res.splitlines():
        if "Function" in ln:
            fn = ln
        elif "REG:" in ln and fn and any(t in fn for t in ("k2_gemm_topk", "k1_stream_kernelILi1ELi3ELi4ELi1ELi2E")):
            assert "LOCAL:0" in ln, f"local memory in {fn}: {ln}"
            stack = int(re.search(r"STACK:(\d+)", ln).group(1))
            assert stack == 0, f"spill stack in {fn}: {ln}"


def test_adjacent_searches_each_distinct_filter_once(lib, monkeypatch):
    """Host logic of the graph-retriever fan-out, no GPU: edges that resolve to the same filter (an edge on a key the
    caller's filter already pins, a repeated edge) become ONE query of the batched search, an edge that contradicts
    the caller's filter becomes the impossible filter, and the merged result is the best k rows over all edges."""
    from githubrepostorag_b200.types import TextNode
    from githubrepostorag_b200.vector_store import B200VectorStore
    st = B200VectorStore(dim=8)
    st.add([TextNode(text=f"t{i}", metadata={"namespace": "ns", "repo": "a" if i < 3 else "b"}, id_=f"n{i}",
                     embedding=[float(i + 1)] * 8) for i in range(6)])
    seen = {}

    def fake_search(queries, k, filters, algo=0, vectors=False):
        assert vectors                      # the fan-out takes the hits' vectors from the search's own result copy
        seen["filters"] = [dict(f) for f in filters]
        seen["queries"] = np.asarray(queries)
        b = len(filters)
        scores = np.full((b, k), -np.inf, np.float32)
        rows = np.full((b, k), -1, np.int64)
        nv = np.zeros(b, np.int32)
        for i, f in enumerate(filters):
            if "__impossible__" in f:
                continue
            hit = [r for r in range(6) if all(str(st._rows.meta(r).get(kk)) == str(v) for kk, v in f.items())][:k]
            for j, r in enumerate(hit):
                scores[i, j], rows[i, j] = 1.0 - 0.1 * r, r
            nv[i] = len(hit)
        return scores, rows, nv, np.zeros((b, k, 8), np.float32)

    class FakeShard:
        row_offset = 0

        def fetch_rows(self, rows):
            return np.zeros((len(rows), 8), np.float32)

        def search_host_any_of(self, q, filt, k, metric="cosine", vectors=False):
            # the union search: rows passing ANY alternative (codes: -1 any, -2 impossible), best k
            seen["alternatives"] = np.asarray(filt).copy()
            codes = st.dictionary.encode_rows([st._rows.meta(r) for r in range(6)])
            hit = [r for r in range(6) if any(all(w == -1 or (w >= 0 and codes[m, r] == w) for m, w in enumerate(alt))
                                              for alt in filt)][:k]
            rows = np.full(k, -1, np.int64)
            rows[:len(hit)] = hit
            return np.zeros(k, np.float32), rows, len(hit), np.zeros((k, 8), np.float32)

    monkeypatch.setattr(st, "_ensure_built", lambda: None)
    monkeypatch.setattr(st, "_search", fake_search)
    monkeypatch.setattr(st, "_shard", FakeShard(), raising=False)
    edges = [("namespace", "ns"), ("repo", "a"), ("repo", "a"), ("namespace", "other"), ("repo", "b")]
    # coded-equality edges leave as ONE union search: four distinct alternatives, the contradicting one impossible
    out = st.adjacent(edges, [0.5] * 8, k=4, filter={"namespace": "ns"})
    alts = seen["alternatives"]
    assert alts.shape == (4, st.dictionary.n_keys) and sum((a == -2).any() for a in alts) == 1
    assert [c.id for c in out] == ["n0", "n1", "n2", "n3"] and "filters" not in seen
    # the per-edge batch (what an edge on a list-valued key, or more than 32 edges, falls back to)
    monkeypatch.setattr(st, "MAX_ALTERNATIVES", 0, raising=False)
    out = st.adjacent(edges, [0.5] * 8, k=4, filter={"namespace": "ns"})
    flt = seen["filters"]
    assert len(flt) == 4                                            # 5 edges: one repeated, none merged wrongly
    assert {"namespace": "ns"} in flt and {"namespace": "ns", "repo": "a"} in flt and {"namespace": "ns", "repo": "b"} in flt
    assert sum("__impossible__" in f for f in flt) == 1            # namespace=other contradicts the caller's filter
    assert seen["queries"].shape == (4, 8)
    assert [c.id for c in out] == ["n0", "n1", "n2", "n3"]          # best four rows over all edges, de-duplicated


def test_per_scope_filter_keys_follow_the_ingest_writer(lib):
    """Every key the reference's writer indexes for a scope, and every key it keeps on all rows, is filterable on
    that scope's table (ingest/src/app/services/vector_write_service.py:28-34,58-64); list-valued keys are served by
    postings, the rest by one code column each, and all of them fit CRB_MAX_KEYS."""
    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.retrievers import GraphRetrieverFactory, TableNames
    from githubrepostorag_b200.vector_store import (ALLOW_FIELDS_BY_SCOPE, KEEP_ALWAYS_KEYS, B200VectorStore,
                                                    filter_keys_for_scope, scope_of_table)
    for scope, allow in ALLOW_FIELDS_BY_SCOPE.items():
        coded, multi = filter_keys_for_scope(scope)
        assert set(coded) | set(multi) == set(allow) | set(KEEP_ALWAYS_KEYS)
        assert set(multi) == set(allow) & {"topics", "imports", "labels"}
        assert len(coded) <= N.MAX_KEYS and len(set(coded)) == len(coded)
    assert {"owner", "symbol"} <= set(filter_keys_for_scope("repo")[0]) and "symbol" in filter_keys_for_scope("chunk")[0]
    assert scope_of_table("embeddings") == "chunk" and scope_of_table("embeddings_repo") == "repo"
    fac = GraphRetrieverFactory(tables=TableNames(repo="embeddings_repo", module="embeddings_module",
                                                  file="embeddings_file", chunk="embeddings"), dim=8)
    st = fac.for_repo().store
    assert st.scope == "repo" and "owner" in st.dictionary.keys and "labels" in st.multi_valued_keys
    assert fac.for_chunk().store.scope == "chunk" and "imports" in fac.for_chunk().store.multi_valued_keys
    # a filter on `owner` is an ordinary coded equality, not the impossible filter
    st.add_embeddings(["r1"], np.ones((1, 8), np.float32), ["t"], [{"owner": "acme", "repo": "x"}])
    eq, extra = st._split_filter({"owner": "acme"})
    assert extra == () and st.dictionary.encode_filter(eq)[st.dictionary.keys.index("owner")] == 0
    assert isinstance(B200VectorStore(dim=8).dictionary.keys, list)


def test_llamaindex_filters_or_nested_in(lib):
    """MetadataFilters -> filter dict: AND leaves stay flat (served by codes); OR and nesting become `__any_of__`
    groups; IN and CONTAINS keep their operators; contradictory equalities are the impossible filter."""
    from githubrepostorag_b200.types import FilterCondition, FilterOperator, MetadataFilter, MetadataFilters
    from githubrepostorag_b200.vector_store import B200VectorStore
    conv = B200VectorStore._filters_to_dict
    F = MetadataFilter
    assert conv(MetadataFilters([F("repo", "a"), F("owner", "o")])) == {"repo": "a", "owner": "o"}
    assert conv(MetadataFilters([F("repo", "a"), F("repo", "b")])) == {"__impossible__": "1"}
    got = conv(MetadataFilters([F("repo", "a"), F("owner", "o")], condition=FilterCondition.OR))
    assert got == {"__any_of__": [[{"repo": "a"}, {"owner": "o"}]]}
    nested = MetadataFilters([F("namespace", "ns"),
                              MetadataFilters([F("topics", "x", FilterOperator.CONTAINS), F("repo", ["a", "b"], FilterOperator.IN)],
                                              condition=FilterCondition.OR)])
    assert conv(nested) == {"namespace": "ns", "__any_of__": [[{"__any_of__": [[{"topics": "x"}, {"repo": ("in", ["a", "b"])}]]}]]}
    st = B200VectorStore(dim=8)
    eq, extra = st._split_filter(conv(nested))
    assert eq == {"namespace": "ns"} and len(extra) == 1 and extra[0][0] == "__any_of__"
    eq, extra = st._split_filter({"topics": ("in", ["b", "a"]), "symbol": "f"})
    assert eq == {"symbol": "f"} and extra == (("topics", ("in", ("a", "b"))),)
    hash(extra)     # the mask cache key


def test_compiled_filters_are_remembered_only_when_every_value_is_known(lib):
    """The store remembers the compiled form of a filter dict (the worker repeats a handful per request); a filter naming
    a value the dictionary has not seen compiles to the impossible filter and is NOT remembered, so a later write that
    introduces the value is found."""
    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.vector_store import B200VectorStore
    st = B200VectorStore(dim=8)
    st.add_embeddings(["a0"], np.ones((1, 8), np.float32), ["t"], [{"repo": "a", "namespace": "ns"}])
    eq, extra, row = st._compile_filter({"repo": "b"})
    assert extra == () and (row == N.FILT_IMPOSSIBLE).any() and not st._filter_cache
    st.add_embeddings(["b0"], np.ones((1, 8), np.float32), ["t"], [{"repo": "b", "namespace": "ns"}])
    hit = st._compile_filter({"repo": "b"})
    pos = st.dictionary.keys.index("repo")
    assert hit[2][pos] == st.dictionary.values[pos]["b"] and st._compile_filter({"repo": "b"}) is hit
    # list-valued keys, IN and OR ride on row masks (extra), unhashable values are compiled without being remembered
    eq, extra, row = st._compile_filter({"repo": "a", "topics": "kafka"})
    assert eq == {"repo": "a"} and extra and row is not None
    n_cached = len(st._filter_cache)
    eq, extra, row = st._compile_filter({"repo": ("in", ["a", "b"])})
    assert eq is None and extra and row is None and len(st._filter_cache) == n_cached
    assert st._compile_filter(None) == (None, (), None) and st._compile_filter({}) == (None, (), None)


def test_write_gate_readers_share_writers_exclude(lib):
    """Searches (readers) run concurrently and re-entrantly; a writer waits for them, excludes them, and is not starved
    by a stream of new readers.  (Event-driven: no outcome depends on how fast a thread gets scheduled.)"""
    import threading
    import time
    from githubrepostorag_b200.vector_store import _WriteGate

    def wait_for(cond, what):
        t0 = time.time()
        while not cond():
            assert time.time() - t0 < 20, f"timed out waiting for {what}"
            time.sleep(0.001)

    gate = _WriteGate()
    log, inside = [], []
    release = threading.Event()

    def reader(i):
        with gate.read():
            with gate.read():                      # re-entrant
                inside.append(i)
                release.wait(20)
        log.append(("r", i))

    def writer():
        with gate.write():
            assert gate._readers == 0
            log.append(("w", len(inside)))
            with gate.write():                     # re-entrant
                pass

    rs = [threading.Thread(target=reader, args=(i,)) for i in range(3)]
    [t.start() for t in rs]
    wait_for(lambda: len(inside) == 3, "three concurrent readers")          # readers share the gate
    w = threading.Thread(target=writer)
    w.start()
    wait_for(lambda: gate._waiting_writers == 1, "the writer to queue")
    assert not any(x[0] == "w" for x in log)                                 # ... and exclude the writer
    release.set()
    [t.join(timeout=20) for t in rs + [w]]
    assert not any(t.is_alive() for t in rs + [w])
    assert ("w", 3) in log and sorted(x for x in log if x[0] == "r") == [("r", 0), ("r", 1), ("r", 2)]
    # a reader arriving while a writer waits queues behind it
    order = []
    hold = threading.Event()
    r1_in = threading.Event()

    def long_reader():
        with gate.read():
            r1_in.set()
            hold.wait(20)
        order.append("r1")

    def late_reader():
        with gate.read():
            order.append("r2")

    def w2():
        with gate.write():
            order.append("w")

    a = threading.Thread(target=long_reader)
    a.start()
    assert r1_in.wait(20)
    b = threading.Thread(target=w2)
    b.start()
    wait_for(lambda: gate._waiting_writers == 1, "the second writer to queue")
    c = threading.Thread(target=late_reader)
    c.start()
    time.sleep(0.05)
    assert order == []                             # the writer waits for r1; the late reader (if it got that far) for the writer
    hold.set()
    [t.join(timeout=20) for t in (a, b, c)]
    assert not any(t.is_alive() for t in (a, b, c)) and order.index("w") < order.index("r2")
    with pytest.raises(RuntimeError):
        with gate.read():
            with gate.write():
                pass

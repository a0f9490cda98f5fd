"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle on the same
seeded inputs.  Bar (BASELINE.json): ids bit-exact except reference ties within 1e-6; scores
within 1e-5 relative (tests/helpers.py states the tolerance)."""
import threading

import numpy as np
import pytest

import oracle
from githubrepostorag_b200 import synthetic as S
from tests.helpers import assert_topk_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def _shard(torch_cuda, rows16, codes=None, row_offset=0, narrow=True):
    from githubrepostorag_b200.index import DeviceShard
    return DeviceShard.from_host(rows16, codes, device=0, row_offset=row_offset, narrow=narrow)


def _check(q, rows16, got_scores, got_rows, n_valid, k, mode, keep=None, what="", row_offset=0):
    e32 = None
    for b in range(len(q)):
        s_ref, r_ref = oracle.query_vectorised(q[b], rows16, k, mode=mode, keep=keep if keep is None or keep.ndim == 1 else keep[b])
        nv = int(n_valid[b])
        assert nv == len(r_ref), f"{what} q{b}: n_valid {nv} vs reference {len(r_ref)}"
        assert np.all(got_rows[b, nv:] == -1) and np.all(np.isneginf(got_scores[b, nv:]))
        if mode == oracle.MODE_COSINE:
            scale = 1.0
        else:
            if e32 is None:
                e32 = np.linalg.norm(rows16.astype(np.float32), axis=1).max()
            scale = float(np.linalg.norm(q[b]) * e32)
        assert_topk_parity(got_rows[b, :nv] - row_offset, got_scores[b, :nv], r_ref, s_ref,
                           lambda row: oracle.score_rows(q[b], rows16[row:row + 1].astype(np.float32), mode)[0],
                           scale=scale, what=f"{what} q{b}")


def _mode_name(mode):
    return "cosine" if mode == oracle.MODE_COSINE else "ip"


# --------------------------------------------------------------------------- golden fixtures
def test_golden_through_store_faces(torch_cuda, golden):
    from githubrepostorag_b200.types import MetadataFilter, MetadataFilters, VectorStoreQuery
    from githubrepostorag_b200.vector_store import B200VectorStore
    for c in golden["cases"]:
        st = B200VectorStore(dim=c["rows16"].shape[1], filter_keys=("namespace", "repo", "branch"),
                             metric=_mode_name(c["mode"]))
        st.add_embeddings(c["ids"], c["rows16"].astype(np.float32), [f"text {i}" for i in c["ids"]], c["metadatas"])
        flt = MetadataFilters(filters=[MetadataFilter(key=k, value=v) for k, v in (c["filters"] or {}).items()])
        res = st.query(VectorStoreQuery(query_embedding=c["query"].tolist(), similarity_top_k=c["k"],
                                        filters=flt if c["filters"] else None))
        e32 = c["rows16"].astype(np.float32)
        scale = 1.0 if c["mode"] == "default" else float(np.linalg.norm(c["query"]) * max(np.linalg.norm(e32, axis=1).max(), 1e-30))
        row = lambda i: int(i.rsplit("-", 1)[1])
        assert_topk_parity([row(i) for i in res.ids], res.similarities, [row(i) for i in c["expected_ids"]],
                           c["expected_sims"], lambda r: oracle.score_rows(c["query"], e32[r:r + 1], c["mode"])[0],
                           scale=scale, what=c["name"])
        assert [n.id_ for n in res.nodes] == res.ids and all(n.text == f"text {n.id_}" for n in res.nodes)
        if c["mode"] == "default":
            triples = st.similarity_search_with_embedding_id_by_vector(c["query"].tolist(), k=c["k"], filter=c["filters"])
            assert [t[2] for t in triples] == res.ids
            for doc, emb, rid in triples:
                assert np.array_equal(np.asarray(emb, np.float32), e32[row(rid)])
                assert doc.id == rid and isinstance(doc.metadata["score"], float)
            scored = st.similarity_search_with_score_id_by_vector(c["query"].tolist(), k=c["k"], filter=c["filters"])
            for (doc, s, rid), cos in zip(scored, res.similarities):
                assert abs(s - oracle.lc_score(cos)) < 1e-7
        st.close()


# --------------------------------------------------------------------------- C1: the reference CPU case
def test_c1_10k_x384_k5_vs_literal_oracle(torch_cuda):
    n, d, k = 10_000, 384, 5
    rows16 = S.corpus(S.BASE_SEEDS["C1"], n, d)
    q = S.queries(S.BASE_SEEDS["C1"], 3, d)
    sh = _shard(torch_cuda, rows16)
    scores, rows, nv = sh.search_host(q, k)
    ids = [S.row_id(i) for i in range(n)]
    e32 = rows16.astype(np.float32)
    for b in range(len(q)):
        sims, out = oracle.query_literal(q[b], e32, ids, k)
        assert_topk_parity(rows[b], scores[b], [int(i[4:]) for i in out], sims,
                           lambda r: oracle.score_rows(q[b], e32[r:r + 1])[0], what=f"C1 q{b}")
    sh.close()


# --------------------------------------------------------------------------- shape sweep
@pytest.mark.parametrize("n,d", [(1, 8), (7, 64), (1000, 8), (3001, 384), (20_000, 768), (5000, 1024),
                                 (4097, 1536), (2000, 4096), (777, 264)])
def test_shape_sweep(torch_cuda, n, d):
    rng = np.random.default_rng(n * 31 + d)
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    sh = _shard(torch_cuda, rows16)
    for mode in (oracle.MODE_COSINE, oracle.MODE_DOT):
        for batch, k in ((1, 1), (1, 10), (2, 5), (3, 32), (4, 33), (5, 100), (9, 128)):
            q = rng.standard_normal((batch, d), dtype=np.float32)
            scores, rows, nv = sh.search_host(q, k, metric=_mode_name(mode))
            _check(q, rows16, scores, rows, nv, k, mode, what=f"n={n} d={d} B={batch} k={k} {mode}")
    sh.close()


def test_ragged_and_degenerate_rows(torch_cuda):
    rng = np.random.default_rng(99)
    rows16 = rng.standard_normal((500, 64), dtype=np.float32).astype(np.float16)
    rows16[10] = 0            # zero-norm row: never returned under cosine
    rows16[20] = rows16[30]   # exact duplicates
    rows16[40] = rows16[30]
    q = np.stack([rows16[30].astype(np.float32), rng.standard_normal(64, dtype=np.float32)])
    sh = _shard(torch_cuda, rows16)
    scores, rows, nv = sh.search_host(q, 3)
    assert sorted(rows[0].tolist()) == [20, 30, 40] and rows[0].tolist() == [40, 30, 20]
    assert scores[0, 0] == scores[0, 1] == scores[0, 2]
    scores, rows, nv = sh.search_host(q, 128)
    assert 10 not in rows[0].tolist()
    _check(q, rows16, scores, rows, nv, 128, oracle.MODE_COSINE, what="degenerate")
    scores, rows, nv = sh.search_host(np.zeros((1, 64), np.float32), 4)   # zero query: scores 0
    assert nv[0] == 4 and np.all(scores[0] == 0)
    sh.close()


def test_empty_shard_and_bad_arguments(torch_cuda):
    sh = _shard(torch_cuda, np.zeros((0, 64), np.float16))
    scores, rows, nv = sh.search_host(np.ones((2, 64), np.float32), 5)
    assert nv.tolist() == [0, 0] and np.all(rows == -1)
    with pytest.raises(RuntimeError, match="k must be"):
        sh.search_host(np.ones((1, 64), np.float32), 129)
    with pytest.raises(ValueError):
        sh.search_host(np.ones((1, 32), np.float32), 5)
    with pytest.raises(ValueError):
        sh.search_host(np.ones((1, 64), np.float32), 5, metric="euclidean")
    sh.close()


# --------------------------------------------------------------------------- filters (K3)
def test_filter_selectivity_sweep(torch_cuda):
    n, d, k = 200_000, 128, 20
    seed = S.BASE_SEEDS["C4"]
    rows16 = S.corpus(seed, n, d)
    codes = S.metadata_codes(seed, n)
    keys = list(S.FILTER_KEYS)
    code_mat = np.stack([codes[kk] for kk in keys]).astype(np.int32)
    code_mat[1, ::97] = -1     # some rows lack `repo`
    sh = _shard(torch_cuda, rows16, code_mat)
    q = S.queries(seed, 2, d)
    ANY = -1
    cases = {
        "all": [ANY, ANY, ANY, ANY],
        "ns0 (~48%)": [0, ANY, ANY, ANY],
        "ns1+dev (~2.4%)": [1, ANY, 1, ANY],
        "repo0+standalone (~1.7%)": [ANY, 0, ANY, 1],
        "repo1500 (~0.004%)": [ANY, 1500, ANY, ANY],
        "ns3+repo1999+dev+standalone (likely 0 rows)": [3, 1999, 1, 1],
        "impossible": [ANY, -2, ANY, ANY],
    }
    for name, want in cases.items():
        keep = np.ones(n, dtype=bool)
        for m, w in enumerate(want):
            if w == -2:
                keep[:] = False
            elif w != ANY:
                keep &= code_mat[m] == w
        filt = np.tile(np.asarray(want, np.int32), (len(q), 1))
        scores, rows, nv = sh.search_host(q, k, filt=filt)
        _check(q, rows16, scores, rows, nv, k, oracle.MODE_COSINE, keep=keep, what=f"filter {name} ({keep.mean():.5f})")
    sh.close()


def test_many_filters_one_query_vector(torch_cuda):
    """graph-retriever adjacent(): one q, one filter per edge -- also > 32 distinct filters
    (two K3 rounds) mixed with unfiltered queries in the same batch."""
    n, d, k = 50_000, 64, 8
    rng = np.random.default_rng(7)
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    code_mat = np.stack([rng.integers(0, 3, n), rng.integers(0, 50, n)]).astype(np.int32)
    sh = _shard(torch_cuda, rows16, code_mat)
    qv = rng.standard_normal(d, dtype=np.float32)
    wants = [[-1, -1]] + [[-1, r] for r in range(40)] + [[1, r] for r in range(5)] + [[-1, -1], [2, -1]]
    q = np.tile(qv, (len(wants), 1))
    filt = np.asarray(wants, np.int32)
    scores, rows, nv = sh.search_host(q, k, filt=filt)
    keep = np.ones((len(wants), n), dtype=bool)
    for i, w in enumerate(wants):
        for m in range(2):
            if w[m] >= 0:
                keep[i] &= code_mat[m] == w[m]
    _check(q, rows16, scores, rows, nv, k, oracle.MODE_COSINE, keep=keep, what="adjacent fan-out")
    sh.close()


@pytest.mark.parametrize("d,k", [(384, 8), (768, 40), (1536, 5), (256, 8), (72, 40)])
def test_skewed_fan_out_balanced_kernel(torch_cuda, d, k):
    """adjacent() fan-out with row sets from a handful of rows to the whole table: the balanced kernel
    (grid split by rows, not by group) must return exactly what the per-group kernel returns, and both
    must match the oracle -- including a filter nothing passes and an impossible one."""
    import os
    torch = torch_cuda
    n = 120_000
    rng = np.random.default_rng(100 + d)
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    repo = (rng.zipf(1.4, n) % 300).astype(np.int32)            # repo 1 holds a large share, most are tiny
    fpath = rng.integers(0, 20_000, n).astype(np.int32)         # ~6 rows per file
    code_mat = np.stack([np.zeros(n, np.int32), repo, fpath])
    sh = _shard(torch, rows16, code_mat)
    qv = rng.standard_normal(d, dtype=np.float32)
    wants = [[0, -1, -1], [-1, 1, -1], [-1, 2, -1], [-1, 250, -1], [0, 1, 77], [-1, -1, 123], [-1, -1, 19_999],
             [-1, 299, 5], [-1, 999, -1], [-2, -1, -1], [-1, 7, -1], [-1, 3, -1]]
    q = np.tile(qv, (len(wants), 1))
    filt = np.asarray(wants, np.int32)
    qd = torch.from_numpy(q).cuda()
    keys_a = torch.empty((len(wants), k), dtype=torch.int64, device="cuda")
    keys_b = torch.empty_like(keys_a)
    os.environ["CODERAG_B200_K1_BALANCED"] = "1"
    try:
        sh.reload_tuning()
        sa, ra, va = [t.clone() for t in sh.search_device(qd, k, filt=filt, keys_out=keys_a)]
        torch.cuda.synchronize()
        os.environ["CODERAG_B200_K1_BALANCED"] = "0"
        sh.reload_tuning()
        sb, rb, vb = [t.clone() for t in sh.search_device(qd, k, filt=filt, keys_out=keys_b)]
        torch.cuda.synchronize()
    finally:
        os.environ.pop("CODERAG_B200_K1_BALANCED", None)
        sh.reload_tuning()
    assert torch.equal(ra, rb) and torch.equal(va, vb) and torch.equal(sa.view(torch.int32), sb.view(torch.int32))
    assert torch.equal(keys_a, keys_b)
    keep = np.ones((len(wants), n), dtype=bool)
    for i, w in enumerate(wants):
        for m in range(3):
            if w[m] == -2:
                keep[i] = False
            elif w[m] >= 0:
                keep[i] &= code_mat[m] == w[m]
    assert va.cpu().numpy()[8] == 0 and va.cpu().numpy()[9] == 0
    _check(q, rows16, sa.cpu().numpy(), ra.cpu().numpy(), va.cpu().numpy(), k, oracle.MODE_COSINE, keep=keep,
           what=f"skewed fan-out d={d}")
    # twice in a row on the same workspace-less path: arrival counters must have reset themselves
    s2, r2, v2 = sh.search_device(qd, k, filt=filt)
    torch.cuda.synchronize()
    assert torch.equal(r2, ra) and torch.equal(s2.view(torch.int32), sa.view(torch.int32))
    sh.close()


@pytest.mark.parametrize("n,d,k", [(100_000, 384, 10), (77_777, 768, 40), (50_001, 1024, 20), (3_000, 64, 5)])
def test_fused_filter_and_score_single_query(torch_cuda, n, d, k):
    """One query under one filter runs the fused filter+score kernel: same rows and score bits as the
    K3 + K1 pair it replaces, and the oracle's answer -- across selectivities from nothing to everything,
    odd row counts (unaligned code columns -> scalar loads), a caller row mask, and the keys output."""
    import os
    torch = torch_cuda
    rng = np.random.default_rng(n + d)
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    code_mat = np.stack([rng.integers(0, 2, n), (rng.zipf(1.5, n) % 40), rng.integers(0, 1000, n)]).astype(np.int32)
    sh = _shard(torch, rows16, code_mat)                    # int8, int8 and int16 columns: what the fused kernel reads
    q = rng.standard_normal((1, d), dtype=np.float32)
    qd = torch.from_numpy(q).cuda()
    extra = rng.random(n) < 0.3
    mask_dev = sh.pack_row_mask(torch.from_numpy(extra).cuda())
    cases = [([0, -1, -1], None), ([-1, 1, -1], None), ([1, 3, -1], None), ([-1, -1, 7], None), ([1, 39, 999], None),
             ([-1, 55, -1], None), ([0, 1, -1], mask_dev), ([-1, -1, -1], mask_dev)]
    for want, mask in cases:
        filt = np.asarray([want], np.int32)
        out = {}
        for fused in ("1", "0"):
            os.environ["CODERAG_B200_K1_FUSED"] = fused
            try:
                sh.reload_tuning()
                keys = torch.empty((1, k), dtype=torch.int64, device="cuda")
                s_, r_, v_ = sh.search_device(qd, k, filt=filt, keys_out=keys, row_masks=None if mask is None else [mask])
                torch.cuda.synchronize()
                out[fused] = (s_.clone(), r_.clone(), v_.clone(), keys.clone())
            finally:
                os.environ.pop("CODERAG_B200_K1_FUSED", None)
                sh.reload_tuning()
        a, b = out["1"], out["0"]
        assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2]), f"rows differ for {want}"
        assert torch.equal(a[0].view(torch.int32), b[0].view(torch.int32)) and torch.equal(a[3], b[3])
        keep = np.ones(n, dtype=bool)
        for m in range(3):
            if want[m] >= 0:
                keep &= code_mat[m] == want[m]
        if mask is not None:
            keep &= extra
        _check(q, rows16, a[0].cpu().numpy(), a[1].cpu().numpy(), a[2].cpu().numpy(), k, oracle.MODE_COSINE, keep=keep,
               what=f"fused n={n} d={d} want={want} mask={mask is not None}")
    # inner product, and twice in a row (the arrival counter resets itself)
    filt = np.asarray([[1, -1, -1]], np.int32)
    keep = code_mat[0] == 1
    for _ in range(2):
        s_, r_, v_ = sh.search_device(qd, k, filt=filt, metric="ip")
        torch.cuda.synchronize()
        _check(q, rows16, s_.cpu().numpy(), r_.cpu().numpy(), v_.cpu().numpy(), k, oracle.MODE_DOT, keep=keep, what="fused ip")
    sh.close()


@pytest.mark.parametrize("d,knob", [(384, "CODERAG_B200_K1_E3"), (768, "CODERAG_B200_K1_EX")])
def test_exact_width_instantiations_return_the_same_bits(torch_cuda, d, knob):
    """dim 384 and dim 768 run streaming-kernel instantiations without per-chunk range tests; they must return
    exactly what the generic instantiation returns (same canonical summation order), filtered or not, both k widths."""
    import os
    torch = torch_cuda
    n = 90_001
    rng = np.random.default_rng(d)
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    code_mat = (np.arange(n) % 7).astype(np.int32)[None, :].copy()
    sh = _shard(torch, rows16, code_mat)
    qd = torch.from_numpy(rng.standard_normal((1, d), dtype=np.float32)).cuda()
    for k in (10, 64):
        for filt in (None, np.asarray([[3]], np.int32)):
            out = {}
            for v in ("1", "0"):
                os.environ[knob] = v
                try:
                    sh.reload_tuning()
                    out[v] = [t.clone() for t in sh.search_device(qd, k, filt=filt, algo=1)]
                    torch.cuda.synchronize()
                finally:
                    os.environ.pop(knob, None)
                    sh.reload_tuning()
            a, b = out["1"], out["0"]
            assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2]) and torch.equal(a[0].view(torch.int32), b[0].view(torch.int32))
    sh.close()


def test_store_adjacent_and_retriever_faces(torch_cuda):
    from githubrepostorag_b200.vector_store import B200VectorStore
    rng = np.random.default_rng(3)
    n, d = 400, 32
    emb = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    metas = [{"namespace": "default", "repo": f"r{i % 5}", "module": f"m{i % 11}", "scope": "chunk"} for i in range(n)]

    class Embedder:
        def embed_query(self, text):
            return emb[int(text)].tolist()

        def embed_documents(self, texts):
            return [emb[int(t)].tolist() for t in texts]

    st = B200VectorStore(Embedder(), dim=d)
    from githubrepostorag_b200.types import Document
    st.add_documents([Document(page_content=str(i), metadata=metas[i]) for i in range(n)], ids=[S.row_id(i) for i in range(n)])
    docs = st.as_retriever(search_kwargs={"k": 4}).invoke("17", filter={"namespace": "default"})
    assert docs[0].id == S.row_id(17) and abs(docs[0].metadata["score"] - 1.0) < 1e-6 and len(docs) == 4
    nws = st.as_retriever(similarity_top_k=1).retrieve("health check".replace("health check", "5"))
    assert len(nws) == 1 and nws[0].node.id_ == S.row_id(5) and abs(nws[0].score - 1.0) < 1e-6
    qvec, first = st.search_with_embedding("17", k=3, filter={"namespace": "default"})
    assert first[0].id == S.row_id(17) and np.allclose(first[0].embedding, emb[17])
    edges = {("repo", "r2"), ("module", "m6"), ("repo", "r2")}
    adj = st.adjacent(edges, qvec, k=6, filter={"namespace": "default"})
    keep = np.array([(m["repo"] == "r2") or (m["module"] == "m6") for m in metas])
    s_ref, r_ref = oracle.query_vectorised(np.asarray(qvec, np.float32), emb.astype(np.float16), 6, keep=keep)
    assert [c.id for c in adj] == [S.row_id(int(r)) for r in r_ref]
    assert [c.id for c in st.get([S.row_id(3), "missing", S.row_id(3)])] == [S.row_id(3)]
    assert st.similarity_search("17", k=2, filter={"repo": "nope"}) == []
    st.close()


# --------------------------------------------------------------------------- ingest-side conversion
def test_encode_rows_matches_numpy_fp16_rounding(torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(21)
    src = (rng.standard_normal((3000, 384), dtype=np.float32) * 3).astype(np.float32)
    src[5] = 0
    src[6, :4] = [65504.0, 1e-8, -6e-8, 70000.0]   # max half, underflow, subnormal, overflow -> inf
    sh = _shard(torch_cuda, src)
    got16 = sh.rows.cpu().numpy()
    with np.errstate(over="ignore"):
        want16 = src.astype(np.float16)
    assert np.array_equal(got16.view(np.uint16), want16.view(np.uint16))
    inv = sh.inv_norm.cpu().numpy()
    with np.errstate(divide="ignore", over="ignore"):
        ref = 1.0 / np.linalg.norm(want16.astype(np.float64), axis=1)
    ok = np.isfinite(ref)
    assert np.allclose(inv[ok], ref[ok], rtol=2e-6) and inv[5] == 0
    sh.close()


# --------------------------------------------------------------------------- sharding + merge
def test_sharded_equals_unsharded_bit_for_bit(torch_cuda):
    torch = torch_cuda
    from githubrepostorag_b200.index import merge_keys
    n, d, k, b = 30_000, 384, 10, 3
    rows16 = S.corpus(77, n, d)
    q = S.queries(77, b, d)
    whole = _shard(torch_cuda, rows16)
    qd = torch.from_numpy(q).cuda()
    keys_whole = torch.empty((b, k), dtype=torch.int64, device="cuda")
    s0, r0, v0 = whole.search_device(qd, k, keys_out=keys_whole)
    cuts = [0, 9_999, 10_000, 23_456, n]
    keys = torch.empty((len(cuts) - 1, b, k), dtype=torch.int64, device="cuda")
    shards = []
    for i in range(len(cuts) - 1):
        sh = _shard(torch_cuda, rows16[cuts[i]:cuts[i + 1]], row_offset=cuts[i])
        sh.search_device(qd, k, keys_out=keys[i])
        shards.append(sh)
    s1, r1, v1, k1 = merge_keys(keys, qd, k)
    torch.cuda.synchronize()
    assert torch.equal(r0, r1) and torch.equal(v0, v1) and torch.equal(keys_whole, k1)
    assert torch.equal(s0.view(torch.int32), s1.view(torch.int32))
    _check(q, rows16, s1.cpu().numpy(), r1.cpu().numpy(), v1.cpu().numpy(), k, oracle.MODE_COSINE, what="sharded")
    for sh in shards + [whole]:
        sh.close()


def test_device_path_equals_host_path_and_is_repeatable(torch_cuda):
    torch = torch_cuda
    n, d, k = 40_000, 768, 10
    rows16 = S.corpus(5, n, d)
    sh = _shard(torch_cuda, rows16)
    q = S.queries(5, 4, d)
    hs, hr, hv = sh.search_host(q, k)
    qd = torch.from_numpy(q).cuda()
    ws = sh.new_workspace(4, k, 0)
    for _ in range(3):
        ds, dr, dv = sh.search_device(qd, k, workspace=ws)
        torch.cuda.synchronize()
        assert np.array_equal(dr.cpu().numpy(), hr) and np.array_equal(ds.cpu().numpy(), hs)
    sh.close()


def test_concurrent_callers_share_one_store(torch_cuda):
    """worker.py runs up to 10 jobs on a thread pool against one store."""
    n, d, k = 60_000, 384, 5
    rows16 = S.corpus(9, n, d)
    sh = _shard(torch_cuda, rows16)
    qs = S.queries(9, 16, d)
    want = [sh.search_host(qs[i:i + 1], k) for i in range(16)]
    errors = []

    def work(i):
        try:
            for _ in range(5):
                s, r, v = sh.search_host(qs[i:i + 1], k)
                assert np.array_equal(r, want[i][1]) and np.array_equal(s, want[i][0])
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(16)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors
    sh.close()


# --------------------------------------------------------------------------- C2 at full size
def test_c2_1m_x768_k10_full_size(torch_cuda):
    n, d, k = 1_000_000, 768, 10
    seed = S.BASE_SEEDS["C2"]
    rows16 = S.corpus(seed, n, d)
    sh = _shard(torch_cuda, rows16)
    q = S.queries(seed, 2, d)
    scores, rows, nv = sh.search_host(q[:1], k)
    _check(q[:1], rows16, scores, rows, nv, k, oracle.MODE_COSINE, what="C2")
    planted = np.array([3, 262_144, 999_999])
    pq = S.planted_queries(seed, rows16[planted])
    scores, rows, nv = sh.search_host(pq, k)
    assert rows[:, 0].tolist() == planted.tolist()      # size-independent property: planted top-1
    assert np.all(scores[:, 0] > 0.99)
    sh.close()


def test_replayed_host_call_equals_the_enqueued_one(torch_cuda, monkeypatch):
    """The one-query unfiltered host call is recorded as a graph on its second occurrence and replayed afterwards: same
    bits as the enqueued form and the device path, across a change of k, of the metric, an append (re-bind), a
    tombstone, filtered calls in between -- and the launch count per call does not change."""
    from githubrepostorag_b200 import _native as N
    torch = torch_cuda
    n, d = 30_011, 768
    rows16 = S.corpus(21, n, d)
    sh = _shard(torch, rows16, (np.arange(n) % 5).astype(np.int32)[None, :].copy())
    qs = S.queries(21, 6, d)

    def both(q, k, metric="cosine", filt=None):
        out = []
        for g in ("1", "0"):
            monkeypatch.setenv("CODERAG_B200_HOST_GRAPH", g)
            sh.reload_tuning()
            res = []
            for _ in range(3):                      # enqueued, recorded + replayed, replayed
                before = N.launch_count()
                r = sh.search_host(q, k, metric=metric, filt=filt)
                res.append(([a.copy() for a in r], N.launch_count() - before))
            out.append(res)
        monkeypatch.delenv("CODERAG_B200_HOST_GRAPH")
        sh.reload_tuning()
        base = out[1][0][0]
        for res in out:
            for r, launches in res:
                assert launches == out[1][0][1]
                assert np.array_equal(r[1], base[1]) and np.array_equal(r[2], base[2]) and np.array_equal(r[0].view(np.int32), base[0].view(np.int32))
        return base

    for i, (k, metric) in enumerate(((10, "cosine"), (64, "cosine"), (10, "ip"), (10, "cosine"))):
        s_, r_, v_ = both(qs[i:i + 1], k, metric)
        ds, dr, dv = sh.search_device(torch.from_numpy(qs[i:i + 1]).cuda(), k, metric=metric)
        torch.cuda.synchronize()
        assert np.array_equal(r_, dr.cpu().numpy()) and np.array_equal(s_.view(np.int32), ds.cpu().numpy().view(np.int32))
    # the graph stays valid across filtered calls, dies with a re-bind, and sees a tombstone (same buffers, new contents)
    for _ in range(3):
        sh.search_host(qs[:1], 10)
    top = int(sh.search_host(qs[:1], 10)[1][0, 0])
    sh.search_host(qs[:1], 10, filt=np.asarray([[3]], np.int32))
    sh.tombstone([top])
    assert top not in sh.search_host(qs[:1], 10)[1][0].tolist()
    more = (qs[0] * 2).astype(np.float16)[None, :]
    sh.append(more, np.asarray([[1]], np.int32))
    for _ in range(3):
        assert int(sh.search_host(qs[:1], 10)[1][0, 0]) == n          # the appended row is the best hit now
    live = np.ones(n + 1, bool)
    live[top] = False
    _check(qs[:1], np.concatenate([rows16, more]), *sh.search_host(qs[:1], 10), 10, oracle.MODE_COSINE, keep=live, what="after append")
    sh.close()

"""CPU tests of the oracle: golden vectors, literal == vectorised, tie and edge semantics."""
import numpy as np
import pytest

import oracle
from tests.helpers import assert_topk_parity


def _row(i):
    return int(i.rsplit("-", 1)[1])


def test_golden_cases_literal(golden):
    assert len(golden["cases"]) >= 20
    for c in golden["cases"]:
        keep = oracle.prefilter_mask(c["metadatas"], c["filters"])
        sims, ids = oracle.query_literal(c["query"], c["rows16"].astype(np.float32), c["ids"], c["k"],
                                         mode=c["mode"], keep=keep)
        assert ids == c["expected_ids"], c["name"]
        np.testing.assert_allclose(sims, c["expected_sims"], rtol=0, atol=0, err_msg=c["name"])


def test_golden_hand_checked_present(golden):
    hand = [c for c in golden["cases"] if c["hand_checked"]]
    assert len(hand) >= 10
    names = {c["name"] for c in golden["cases"]}
    for must in ("zero_norm_row", "k_exceeds_rows", "filter_unknown_key", "filter_unknown_value",
                 "duplicate_vectors_boundary", "filter_exactly_k", "k_equals_1"):
        assert must in names


def test_vectorised_matches_literal_on_golden(golden):
    for c in golden["cases"]:
        keep = oracle.prefilter_mask(c["metadatas"], c["filters"])
        e32 = c["rows16"].astype(np.float32)
        s, r = oracle.query_vectorised(c["query"], c["rows16"], c["k"], mode=c["mode"], keep=keep, block_rows=17)
        ref_rows = [_row(i) for i in c["expected_ids"]]
        scale = 1.0 if c["mode"] == "default" else float(np.linalg.norm(c["query"]) * np.linalg.norm(e32, axis=1).max())
        assert_topk_parity(r, s, ref_rows, c["expected_sims"],
                           lambda row: oracle.score_rows(c["query"], e32[row:row + 1], c["mode"])[0],
                           scale=scale, what=c["name"])


@pytest.mark.parametrize("n,d,k", [(1000, 384, 5), (3000, 64, 10), (513, 768, 32)])
def test_vectorised_matches_literal_random(n, d, k):
    rng = np.random.default_rng(n + d)
    e16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    e32 = e16.astype(np.float32)
    q = rng.standard_normal(d, dtype=np.float32)
    ids = [f"row-{i:09d}" for i in range(n)]
    keep = rng.random(n) < 0.6
    for mode in (oracle.MODE_COSINE, oracle.MODE_DOT):
        sims, out = oracle.query_literal(q, e32, ids, k, mode=mode, keep=keep)
        s, r = oracle.query_vectorised(q, e16, k, mode=mode, keep=keep, block_rows=256)
        scale = 1.0 if mode == oracle.MODE_COSINE else float(np.linalg.norm(q) * np.linalg.norm(e32, axis=1).max())
        assert_topk_parity(r, s, [_row(i) for i in out], sims,
                           lambda row: oracle.score_rows(q, e32[row:row + 1], mode)[0], scale=scale)


def test_heap_keeps_larger_ids_on_exact_ties():
    e = np.zeros((6, 8), np.float32)
    e[:, 0] = 1.0
    q = e[0]
    ids = [f"row-{i:09d}" for i in range(6)]
    sims, out = oracle.query_literal(q, e, ids, 3)
    assert sorted(out) == ["row-000000003", "row-000000004", "row-000000005"]
    s, r = oracle.query_vectorised(q, e.astype(np.float16), 3)
    assert list(r) == [5, 4, 3]


def test_fp32_oracle_is_close_to_fp64():
    rng = np.random.default_rng(5)
    e16 = rng.standard_normal((2000, 768), dtype=np.float32).astype(np.float16)
    q = rng.standard_normal(768, dtype=np.float32)
    s32 = oracle.score_rows(q, e16.astype(np.float32))
    s64 = oracle.score_rows_f64(q, e16)
    assert np.max(np.abs(s32 - s64)) < 2e-7


def test_prefilter_semantics():
    md = [{"a": "1", "b": "x"}, {"a": "1"}, {"a": "2", "b": "x"}, {}]
    assert oracle.prefilter_mask(md, None).tolist() == [True] * 4
    assert oracle.prefilter_mask(md, {"a": "1"}).tolist() == [True, True, False, False]
    assert oracle.prefilter_mask(md, {"a": "1", "b": "x"}).tolist() == [True, False, False, False]
    assert oracle.prefilter_mask(md, {"c": "1"}).tolist() == [False] * 4
    ids = ["p", "q", "r", "s"]
    assert oracle.prefilter_mask(md, None, ids=ids, node_ids=["q", "s"]).tolist() == [False, True, False, True]


def test_lc_score_convention():
    assert oracle.lc_score(1.0) == 1.0 and oracle.lc_score(-1.0) == 0.0 and oracle.lc_score(0.0) == 0.5


@pytest.mark.parametrize("n,d,k", [(2000, 384, 10), (777, 768, 5), (4096, 64, 40)])
def test_oracle_agrees_with_scikit_learn_brute_force(n, d, k):
    """An independent second implementation (scikit-learn's cosine_similarity / brute-force NearestNeighbors)
    ranks the same rows as the oracle restatement and scores them within 1e-6.  It does not pin the oracle to
    the reference (nothing can: the reference holds no vector), but the goldens are no longer self-confirmed."""
    sk_pair = pytest.importorskip("sklearn.metrics.pairwise")
    sk_nn = pytest.importorskip("sklearn.neighbors")
    rng = np.random.default_rng(7 * n + d)
    e16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    e64 = e16.astype(np.float64)
    q = rng.standard_normal(d, dtype=np.float32)
    ids = [f"row-{i:09d}" for i in range(n)]
    sims, out = oracle.query_literal(q, e16.astype(np.float32), ids, k)
    s_vec, r_vec = oracle.query_vectorised(q, e16, k)
    cos = sk_pair.cosine_similarity(q[None, :].astype(np.float64), e64)[0]
    order = np.lexsort((-np.arange(n), -cos))[:k]
    assert [_row(i) for i in out] == order.tolist() == r_vec.tolist()
    np.testing.assert_allclose(sims, cos[order], rtol=0, atol=1e-6)
    nn = sk_nn.NearestNeighbors(n_neighbors=k, algorithm="brute", metric="cosine").fit(e64)
    dist, idx = nn.kneighbors(q[None, :].astype(np.float64))
    assert idx[0].tolist() == order.tolist()
    np.testing.assert_allclose(1.0 - dist[0], sims, rtol=0, atol=1e-6)
    # inner product against a plain float64 matrix product
    sims_ip, out_ip = oracle.query_literal(q, e16.astype(np.float32), ids, k, mode=oracle.MODE_DOT)
    dots = e64 @ q.astype(np.float64)
    assert [_row(i) for i in out_ip] == np.lexsort((-np.arange(n), -dots))[:k].tolist()


def test_prefilter_or_in_contains_semantics():
    """AND / OR of EQ, IN and CONTAINS leaves (upstream `_build_metadata_filter_fn`), rows lacking a key never match."""
    mds = [{"repo": "a", "owner": "acme", "topics": ["x", "y"]}, {"repo": "b", "owner": "acme"},
           {"repo": "c", "topics": ["y"]}, {"owner": "zed", "topics": []}, {}]
    m = lambda f, **kw: oracle.prefilter_mask(mds, f, **kw).tolist()   # noqa: E731
    assert m({"owner": "acme"}) == [True, True, False, False, False]
    assert m({"owner": "acme", "repo": "b"}) == [False, True, False, False, False]
    assert m({"owner": "acme", "repo": "c"}, condition="or") == [True, True, True, False, False]
    assert m({"topics": "y"}) == [True, False, True, False, False]                       # CONTAINS on a list
    assert m({"topics": ("in", ["x", "q"])}) == [True, False, False, False, False]      # shares a member
    assert m({"repo": ("in", ["a", "c", "zz"])}) == [True, False, True, False, False]
    assert m({"__any_of__": [[{"repo": "a"}, {"owner": "zed"}]]}) == [True, False, False, True, False]
    assert m({"owner": "acme", "__any_of__": [[{"repo": "b"}, {"topics": "x"}]]}) == [True, True, False, False, False]
    assert m({"nope": "1"}) == [False] * 5


@pytest.mark.parametrize("mode", [oracle.MODE_COSINE, oracle.MODE_DOT])
def test_blockwise_multi_query_form_equals_the_vectorised_form(mode):
    """`query_blocks` (what the full-size GPU tests stream 10M-row corpora through) ranks like `query_vectorised`."""
    rng = np.random.default_rng(99)
    n, d, k = 5000, 96, 12
    e16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    e16[100] = 0
    e16[200:230] = e16[200]                         # exact ties across a block boundary
    Q = rng.standard_normal((5, d), dtype=np.float32)
    Q[1] = e16[200].astype(np.float32)
    keep = rng.random(n) < 0.7
    keep[200:230] = True
    for kp in (None, keep):
        blocks = [(lo, e16[lo:lo + 211], None if kp is None else kp[lo:lo + 211]) for lo in range(0, n, 211)]
        got = oracle.query_blocks(Q, blocks, k, mode=mode)
        for b in range(len(Q)):
            s, r = oracle.query_vectorised(Q[b], e16, k, mode=mode, keep=kp)
            assert got[b][1].tolist() == r.tolist()
            np.testing.assert_allclose(got[b][0], s, rtol=2e-6, atol=1e-6)

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

"""Generate tests/golden/retrieval_golden.json from the oracle (run from the repo root:
``python tests/golden/make_golden.py``).

The reference has no golden vectors for retrieval ("parity unpinned"), so
these are ours: hand-derivable cases (expected ids written down by hand and asserted against
the oracle here) plus seeded cases whose expected output is the literal oracle's
(``oracle.query_literal``: per-row loop + heap, the restated LlamaIndex algorithm).  Corpus
values are fp16-exact and stored as base64 float16 so every consumer sees identical bits.
"""
from __future__ import annotations

import base64
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402

OUT = Path(__file__).resolve().parent / "retrieval_golden.json"


def b64(a: np.ndarray) -> str:
    return base64.b64encode(np.ascontiguousarray(a).tobytes()).decode("ascii")


CORPORA = {}


def case(name, rows16, query, k, mode="default", metadatas=None, filters=None, hand_ids=None, note="", corpus=None):
    rows16 = np.asarray(rows16, dtype=np.float16)
    corpus = corpus or name
    if corpus not in CORPORA:
        CORPORA[corpus] = {"n": len(rows16), "dim": int(rows16.shape[1]), "rows_f16_b64": b64(rows16)}
    n = len(rows16)
    ids = [f"row-{i:09d}" for i in range(n)]
    metadatas = metadatas or [{} for _ in range(n)]
    keep = oracle.prefilter_mask(metadatas, filters)
    sims, out_ids = oracle.query_literal(np.asarray(query, np.float32), rows16.astype(np.float32), ids, k,
                                         mode=mode, keep=keep)
    if hand_ids is not None:
        assert out_ids == hand_ids, (name, out_ids, hand_ids)
    return {
        "name": name, "note": note, "corpus": corpus, "k": k, "mode": mode,
        "query_f32_b64": b64(np.asarray(query, np.float32)),
        "metadatas": metadatas if filters else None, "filters": filters,
        "expected_ids": out_ids, "expected_sims": sims,
        "hand_checked": hand_ids is not None,
    }


def main():
    cases = []
    e = np.eye(8, dtype=np.float32)
    q = e[0] * 2.0
    # --- hand cases (d = 8) ---
    rows = [e[0], -e[0], e[1], e[0] + e[1], 3 * e[0] + 4 * e[1]]
    # cos: 1, -1, 0, 0.7071, 0.6
    cases.append(case("axis_aligned", rows, q, 3, hand_ids=["row-000000000", "row-000000003", "row-000000004"],
                      note="identical / negated / orthogonal / 45deg / 3-4-5 rows"))
    cases.append(case("k_equals_1", rows, q, 1, hand_ids=["row-000000000"]))
    cases.append(case("k_exceeds_rows", rows, q, 10,
                      hand_ids=["row-000000000", "row-000000003", "row-000000004", "row-000000002", "row-000000001"],
                      note="fewer than k survivors returns the short list"))
    cases.append(case("inner_product", rows, q, 2, mode="dot_product",
                      hand_ids=["row-000000004", "row-000000003"], note="dots 2,-2,0,2,6: row 4, then rows 0 and 3 tie at 2 exactly and the heap keeps the larger id"))
    rows_z = [e[0], np.zeros(8), e[1], 0.5 * e[0] + 0.5 * e[2]]
    cases.append(case("zero_norm_row", rows_z, q, 4,
                      hand_ids=["row-000000000", "row-000000003", "row-000000002"],
                      note="zero row is never returned under cosine (documented deviation)"))
    dup = [e[1], e[0] + e[2], e[0] + e[2], e[0] + e[2], e[3]]
    cases.append(case("duplicate_vectors_boundary", dup, q, 2,
                      hand_ids=None, note="three bit-identical rows tie at the k-th boundary: the heap keeps the two LARGER ids"))
    md = [{"namespace": "default", "repo": "a"}, {"namespace": "default", "repo": "b"},
          {"namespace": "acme", "repo": "a"}, {"namespace": "default", "repo": "a", "branch": "dev"},
          {"repo": "a"}]
    rows_f = [e[0] + 0.1 * e[1], e[0], e[0] + 0.2 * e[1], e[0] + 0.3 * e[1], e[0] + 0.05 * e[1]]
    cases.append(case("filter_and", rows_f, q, 3, metadatas=md, filters={"namespace": "default", "repo": "a"},
                      hand_ids=["row-000000000", "row-000000003"], note="row 4 lacks `namespace` -> excluded"))
    cases.append(case("filter_exactly_k", rows_f, q, 2, metadatas=md, filters={"namespace": "default", "repo": "a"},
                      hand_ids=["row-000000000", "row-000000003"]))
    cases.append(case("filter_unknown_value", rows_f, q, 3, metadatas=md, filters={"repo": "zzz"}, hand_ids=[]))
    cases.append(case("filter_unknown_key", rows_f, q, 3, metadatas=md, filters={"owner": "x"}, hand_ids=[]))
    cases.append(case("filter_single_key", rows_f, q, 5, metadatas=md, filters={"repo": "a"},
                      hand_ids=["row-000000004", "row-000000000", "row-000000002", "row-000000003"]))
    # --- seeded cases ---
    for name, n, d, k, seed in (("seeded_64x8", 64, 8, 5, 11), ("seeded_48x384", 48, 384, 5, 12),
                                ("seeded_96x768", 96, 768, 10, 13)):
        rng = np.random.default_rng(seed)
        r16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
        qq = rng.standard_normal(d, dtype=np.float32)
        cases.append(case(name, r16, qq, k, corpus=name))
        cases.append(case(name + "_ip", r16, qq, k, mode="dot_product", corpus=name))
        mds = [{"namespace": ["default", "acme"][int(rng.integers(2))], "repo": f"r{int(rng.integers(4))}"} for _ in range(n)]
        cases.append(case(name + "_filtered", r16, qq, k, metadatas=mds, filters={"namespace": "acme", "repo": "r1"}, corpus=name))
    OUT.write_text(json.dumps({"generator": "tests/golden/make_golden.py", "corpora": CORPORA, "cases": cases},
                              separators=(",", ":")))
    print(f"wrote {len(cases)} cases to {OUT}")


if __name__ == "__main__":
    main()

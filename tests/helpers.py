"""Test-side helpers: golden loading and the tie-aware parity comparator."""
from __future__ import annotations

import base64
import json
from pathlib import Path
from typing import Callable, Dict, List, Sequence

import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden" / "retrieval_golden.json"

# Parity bar (BASELINE.json north_star): ids bit-exact except where reference scores tie
# within 1e-6; scores within 1e-5 relative.  The absolute floor covers scores near zero,
# where cancellation makes any fp32 evaluation (the oracle's included) relative-unstable:
# it is expressed in units of ||q||*||e|| (1 for cosine).
SCORE_RTOL = 1e-5
SCORE_ATOL_UNIT = 2e-7
TIE_EPS = 1e-6


def load_golden() -> Dict:
    raw = json.loads(GOLDEN.read_text())
    corpora = {}
    for name, c in raw["corpora"].items():
        rows = np.frombuffer(base64.b64decode(c["rows_f16_b64"]), dtype=np.float16).reshape(c["n"], c["dim"])
        corpora[name] = rows
    cases = []
    for c in raw["cases"]:
        c = dict(c)
        c["rows16"] = corpora[c["corpus"]]
        c["query"] = np.frombuffer(base64.b64decode(c["query_f32_b64"]), dtype=np.float32)
        n = len(c["rows16"])
        c["ids"] = [f"row-{i:09d}" for i in range(n)]
        c["metadatas"] = c["metadatas"] or [{} for _ in range(n)]
        cases.append(c)
    return {"cases": cases, "corpora": corpora}


def assert_topk_parity(got_rows: Sequence[int], got_scores: Sequence[float],
                       ref_rows: Sequence[int], ref_scores: Sequence[float],
                       ref_score_of: Callable[[int], float], *, scale: float = 1.0, what: str = "") -> None:
    """ids must match position by position, except where the reference scores involved tie
    within TIE_EPS (then any row whose *reference* score ties with the reference's score at
    that position is accepted); scores within SCORE_RTOL relative (+ tiny absolute floor)."""
    assert len(got_rows) == len(ref_rows), f"{what}: {len(got_rows)} hits vs reference {len(ref_rows)}"
    assert len(set(got_rows)) == len(got_rows), f"{what}: duplicate rows returned {list(got_rows)}"
    for i, (gr, gs, rr, rs) in enumerate(zip(got_rows, got_scores, ref_rows, ref_scores)):
        ref_of_got = float(rs)
        if int(gr) != int(rr):
            ref_of_got = float(ref_score_of(int(gr)))
            assert abs(ref_of_got - float(rs)) <= TIE_EPS * max(scale, 1.0), (
                f"{what}: position {i}: row {gr} (reference score {ref_of_got!r}) returned where the "
                f"reference has row {rr} (score {rs!r}); not tied within {TIE_EPS}")
        tol = SCORE_RTOL * abs(ref_of_got) + SCORE_ATOL_UNIT * scale
        assert abs(float(gs) - ref_of_got) <= tol, (
            f"{what}: position {i}: row {gr} score {gs!r} vs reference {ref_of_got!r} (tol {tol:.3e})")
    assert all(float(got_scores[i]) >= float(got_scores[i + 1]) for i in range(len(got_scores) - 1)), \
        f"{what}: scores not descending"

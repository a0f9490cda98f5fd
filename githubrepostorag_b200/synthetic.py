"""Seeded synthetic corpora, queries and metadata of the BASELINE.json shapes.

Rows are generated in blocks of 2**18 with ``numpy.random.default_rng([base_seed, block])`` so
any shard on any host regenerates identical rows without holding the whole corpus.  The fp16
cast of the block *is* the corpus; rows are not pre-normalised.  Metadata follows the
vocabulary the reference writes at ``ingest/src/app/ingest_controller.py:164-188`` and
``ingest/src/app/services/transform_service.py:112-127``.
"""
from __future__ import annotations

from typing import Dict, Iterator, List, Tuple

import numpy as np

BLOCK_ROWS = 1 << 18

BASE_SEEDS = {"C1": 1001, "C2": 1002, "C3": 1003, "C4": 1004, "C5": 1005}

NAMESPACES = ["default", "acme", "labs", "infra"]
BRANCHES = ["main", "dev"]
COMPONENT_KINDS = ["service", "standalone"]
N_REPOS = 2000
FILTER_KEYS = ("namespace", "repo", "branch", "component_kind")


def row_id(i: int) -> str:
    return f"row-{i:09d}"


def corpus_block(base_seed: int, block: int, dim: int, rows: int = BLOCK_ROWS) -> np.ndarray:
    """fp16 rows of one block (``rows`` < BLOCK_ROWS only for the last block)."""
    rng = np.random.default_rng([base_seed, block])
    return rng.standard_normal((rows, dim), dtype=np.float32).astype(np.float16)


def corpus_blocks(base_seed: int, n_rows: int, dim: int, lo: int = 0, hi: int | None = None
                  ) -> Iterator[Tuple[int, np.ndarray]]:
    """Yield ``(first_row, fp16 block)`` covering rows ``[lo, hi)``; blocks are regenerated
    whole and sliced so a shard boundary inside a block is exact."""
    hi = n_rows if hi is None else hi
    b = lo // BLOCK_ROWS
    while b * BLOCK_ROWS < hi:
        first = b * BLOCK_ROWS
        rows = min(BLOCK_ROWS, n_rows - first)
        blk = corpus_block(base_seed, b, dim, rows)
        s = max(lo, first) - first
        e = min(hi, first + rows) - first
        yield first + s, blk[s:e]
        b += 1


def corpus(base_seed: int, n_rows: int, dim: int, lo: int = 0, hi: int | None = None) -> np.ndarray:
    hi = n_rows if hi is None else hi
    out = np.empty((hi - lo, dim), dtype=np.float16)
    for first, blk in corpus_blocks(base_seed, n_rows, dim, lo, hi):
        out[first - lo:first - lo + len(blk)] = blk
    return out


def queries(base_seed: int, batch: int, dim: int) -> np.ndarray:
    """fp32 queries, seed ``base_seed + 1``."""
    rng = np.random.default_rng(base_seed + 1)
    return rng.standard_normal((batch, dim), dtype=np.float32)


def planted_queries(base_seed: int, rows: np.ndarray, noise: float = 0.05) -> np.ndarray:
    """``q = e_j + noise * N(0,1)`` so the top-1 row of each query is known a priori."""
    rng = np.random.default_rng(base_seed + 3)
    r = np.asarray(rows, dtype=np.float32)
    return r + noise * rng.standard_normal(r.shape, dtype=np.float32)


def _zipf_probs(n: int, a: float) -> np.ndarray:
    w = 1.0 / np.arange(1, n + 1, dtype=np.float64) ** a
    return w / w.sum()


def metadata_codes(base_seed: int, n_rows: int, lo: int = 0, hi: int | None = None) -> Dict[str, np.ndarray]:
    """Integer draws per filter key for rows ``[lo, hi)`` (block-seeded with ``base_seed + 2``):
    namespace ~ Zipf over 4, repo ~ Zipf(1.1) over 2000, branch 90/10, component_kind 80/20."""
    hi = n_rows if hi is None else hi
    out = {k: np.empty(hi - lo, dtype=np.int32) for k in FILTER_KEYS}
    p_ns = _zipf_probs(len(NAMESPACES), 1.0)
    p_repo = _zipf_probs(N_REPOS, 1.1)
    b = lo // BLOCK_ROWS
    while b * BLOCK_ROWS < hi:
        first = b * BLOCK_ROWS
        rows = min(BLOCK_ROWS, n_rows - first)
        rng = np.random.default_rng([base_seed + 2, b])
        ns = rng.choice(len(NAMESPACES), size=rows, p=p_ns).astype(np.int32)
        repo = rng.choice(N_REPOS, size=rows, p=p_repo).astype(np.int32)
        br = (rng.random(rows) < 0.10).astype(np.int32)
        ck = (rng.random(rows) < 0.20).astype(np.int32)
        s = max(lo, first) - first
        e = min(hi, first + rows) - first
        d0 = first + s - lo
        for k, col in (("namespace", ns), ("repo", repo), ("branch", br), ("component_kind", ck)):
            out[k][d0:d0 + (e - s)] = col[s:e]
        b += 1
    return out


def vocab() -> Dict[str, List[str]]:
    return {
        "namespace": list(NAMESPACES),
        "repo": [f"repo-{i:04d}" for i in range(N_REPOS)],
        "branch": list(BRANCHES),
        "component_kind": list(COMPONENT_KINDS),
    }


def metadata_dicts(codes: Dict[str, np.ndarray]) -> List[Dict[str, str]]:
    """String metadata rows as the ingest writer stores them (every value a ``str``,
    ``ingest/src/app/services/vector_write_service.py:68-86``)."""
    v = vocab()
    n = len(next(iter(codes.values())))
    return [{k: v[k][int(codes[k][i])] for k in codes} for i in range(n)]


def pick_filter(codes: Dict[str, np.ndarray], keys: Tuple[str, ...], target: float
                ) -> Tuple[Dict[str, str], float]:
    """Scan value combinations of ``keys`` and return the equality filter whose joint
    selectivity is closest to ``target`` together with the achieved selectivity (C4)."""
    v = vocab()
    n = len(codes[keys[0]])
    combo = np.zeros(n, dtype=np.int64)
    radix = []
    for k in keys:
        combo = combo * len(v[k]) + codes[k]
        radix.append(len(v[k]))
    vals, counts = np.unique(combo, return_counts=True)
    j = int(np.argmin(np.abs(counts / n - target)))
    c = int(vals[j])
    picked: Dict[str, str] = {}
    for k, r in zip(reversed(keys), reversed(radix)):
        picked[k] = v[k][c % r]
        c //= r
    return {k: picked[k] for k in keys}, float(counts[j] / n)

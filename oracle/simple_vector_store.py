"""numpy restatement of LlamaIndex's in-memory vector search.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED (see ``oracle/__init__.py``): the upstream modules below are not present
under /root/reference and cannot be imported here; this file restates their published
algorithm.  Upstream symbols restated (llama-index-core >= 0.10.0, the floor pinned at
``environment-worker.yaml:38``):

* ``llama_index/core/base/embeddings/base.py::similarity``            -> :func:`similarity`
* ``llama_index/core/indices/query/embedding_utils.py::get_top_k_embeddings``
                                                                     -> :func:`get_top_k_embeddings`
* ``llama_index/core/vector_stores/simple.py::_build_metadata_filter_fn`` (EQ / IN / CONTAINS, AND / OR)
                                                                     -> :func:`prefilter_mask`
* ``llama_index/core/vector_stores/simple.py::SimpleVectorStore.query`` -> :func:`query_literal`

Reference call sites this anchors on: ``rest_api/src/app/health.py:104-107``
(``index.as_retriever(similarity_top_k=k).retrieve(q)``) and the filter vocabulary built at
``rag_worker/src/worker/services/agent_graph.py:198-239,357-364,532-534`` (flat dict of
string equalities, ANDed).

Two forms are kept:

(i)  :func:`query_literal` -- the per-row loop + ``heapq`` of ``(similarity, id)`` tuples,
     authoritative for tie semantics, practical up to ~100k rows;
(ii) :func:`query_vectorised` -- blocked ``E_blk @ q`` + ``argpartition``; proven equal to (i)
     on small shapes by ``tests/test_oracle.py`` and used as the timed CPU baseline.

dtype: the north star fixes the oracle at **fp32** on the fp16-rounded corpus upcast to fp32
(upstream would build float64 arrays from Python lists).  :func:`score_rows_f64` is used only
to classify near-ties.

Documented deviation: a zero-norm row makes upstream's cosine ``0/0 = nan`` (with a numpy
warning) and the heap order of nan tuples is undefined; here, and in the CUDA path, such rows
are never returned under cosine.
"""
from __future__ import annotations

import heapq
from typing import Any, Dict, List, Mapping, Optional, Sequence, Tuple

import numpy as np

MODE_COSINE = "default"        # VectorStoreQueryMode.DEFAULT
MODE_DOT = "dot_product"       # SimilarityMode.DOT_PRODUCT


def similarity(embedding1: np.ndarray, embedding2: np.ndarray, mode: str = MODE_COSINE):
    """Upstream ``similarity``: cosine is ``dot / (norm * norm)``, computed per row."""
    if mode == "euclidean":
        return -float(np.linalg.norm(np.asarray(embedding1) - np.asarray(embedding2)))
    if mode == MODE_DOT:
        return np.dot(embedding1, embedding2)
    product = np.dot(embedding1, embedding2)
    norm = np.linalg.norm(embedding1) * np.linalg.norm(embedding2)
    return product / norm


def _filter_match(md: Mapping[str, Any], key: str, value: Any) -> bool:
    """Upstream ``_process_filter_match`` for the operators this path uses.  ``value`` is a plain value (EQ; on
    a list-valued metadata entry CONTAINS -- upstream's ``value in metadata_value``) or ``("in", [values])``
    (IN: ``metadata_value in value``).  For list-valued metadata IN means "shares a member": the ingest writer
    shreds such lists into one ``key -> member`` entry each
    (``ingest/src/app/services/vector_write_service.py:152-154``), so each member is matched on its own.
    A row lacking the key never matches."""
    have = md.get(key)
    if have is None:
        return False
    members = list(have) if isinstance(have, (list, tuple, set, frozenset)) else None
    if isinstance(value, tuple) and len(value) == 2 and value[0] == "in":
        allowed = set(value[1])
        return any(m in allowed for m in members) if members is not None else have in allowed
    return value in members if members is not None else have == value


def filter_passes(md: Mapping[str, Any], filters: Optional[Mapping[str, Any]], condition: str = "and") -> bool:
    """One row against a filter dict.  Upstream ``_build_metadata_filter_fn``: every leaf is evaluated, then
    ``all`` (condition AND, the only one the reference emits) or ``any`` (OR).  The key ``"__any_of__"`` holds
    nested groups -- a list of groups, each a list of alternative filter dicts, OR inside a group -- which is
    how LlamaIndex's nested ``MetadataFilters`` / ``FilterCondition.OR`` reach the store under test."""
    if not filters:
        return True
    results = []
    for key, value in filters.items():
        if key == "__any_of__":
            results.extend(any(filter_passes(md, alt) for alt in group) for group in value)
        else:
            results.append(_filter_match(md, key, value))
    return all(results) if condition == "and" else any(results)


def prefilter_mask(
    metadatas: Sequence[Mapping[str, Any]],
    filters: Optional[Mapping[str, Any]] = None,
    *,
    ids: Optional[Sequence[str]] = None,
    node_ids: Optional[Sequence[str]] = None,
    condition: str = "and",
) -> np.ndarray:
    """Upstream ``_build_metadata_filter_fn``: AND (or OR) of EQ / IN / CONTAINS predicates; a row lacking the
    key has ``metadata.get(key) is None`` and is excluded.  ``node_ids`` is the optional allow-set of
    ``VectorStoreQuery.node_ids``."""
    n = len(metadatas)
    keep = np.ones(n, dtype=bool)
    if filters:
        for i, md in enumerate(metadatas):
            keep[i] = filter_passes(md, filters, condition)
    if node_ids is not None:
        assert ids is not None
        allow = set(node_ids)
        for i, rid in enumerate(ids):
            if rid not in allow:
                keep[i] = False
    return keep


def get_top_k_embeddings(
    query_embedding: np.ndarray,
    embeddings: np.ndarray,
    *,
    similarity_top_k: Optional[int] = None,
    embedding_ids: Optional[Sequence[Any]] = None,
    similarity_cutoff: Optional[float] = None,
    mode: str = MODE_COSINE,
) -> Tuple[List[float], List[Any]]:
    """Upstream ``get_top_k_embeddings``: push every row's ``(similarity, id)`` on a min-heap,
    pop when the heap exceeds k (so on an exact tie the *smaller* id tuple is evicted first),
    then ``sorted(heap, key=similarity, reverse=True)``."""
    if embedding_ids is None:
        embedding_ids = list(range(len(embeddings)))
    q = np.asarray(query_embedding)
    heap: List[Tuple[Any, Any]] = []
    for i, emb in enumerate(embeddings):
        if mode == MODE_COSINE and not np.any(emb):
            continue  # documented deviation: zero-norm rows are never returned
        sim = similarity(q, emb, mode)
        if similarity_cutoff is None or sim > similarity_cutoff:
            heapq.heappush(heap, (sim, embedding_ids[i]))
            if similarity_top_k and len(heap) > similarity_top_k:
                heapq.heappop(heap)
    result = sorted(heap, key=lambda t: t[0], reverse=True)
    return [float(s) for s, _ in result], [rid for _, rid in result]


def query_literal(
    query_embedding: np.ndarray,
    embeddings: np.ndarray,
    ids: Sequence[str],
    k: int,
    *,
    mode: str = MODE_COSINE,
    keep: Optional[np.ndarray] = None,
) -> Tuple[List[float], List[str]]:
    """Upstream ``SimpleVectorStore.query``: prefilter, then ``get_top_k_embeddings`` over the
    surviving rows in insertion order.  Returns ``(similarities, ids)`` sorted descending."""
    q = np.asarray(query_embedding, dtype=np.float32)
    emb = np.asarray(embeddings, dtype=np.float32)
    if keep is None:
        rows = range(len(emb))
    else:
        rows = np.flatnonzero(keep)
    sel_emb = [emb[i] for i in rows]
    sel_ids = [ids[i] for i in rows]
    return get_top_k_embeddings(q, sel_emb, similarity_top_k=k, embedding_ids=sel_ids, mode=mode)


def _row_norms(block: np.ndarray) -> np.ndarray:
    return np.sqrt(np.einsum("ij,ij->i", block, block, dtype=np.float32)).astype(np.float32)


def query_vectorised(
    query_embedding: np.ndarray,
    embeddings: np.ndarray,
    k: int,
    *,
    mode: str = MODE_COSINE,
    keep: Optional[np.ndarray] = None,
    block_rows: int = 1 << 16,
    row_offset: int = 0,
) -> Tuple[np.ndarray, np.ndarray]:
    """Blocked form of :func:`query_literal` for 1M-100M rows (the timed CPU baseline).

    ``embeddings`` may be float16 (upcast per block to fp32, which is the corpus the oracle is
    defined on) or float32.  Returns ``(similarities fp32[k'], rows int64[k'])`` with
    ``k' = min(k, survivors)``, ordered by similarity descending and, among equal
    similarities, by row index descending -- the order in which the literal heap retains
    zero-padded ``row-%09d`` ids at the k-th boundary.
    """
    q = np.asarray(query_embedding, dtype=np.float32)
    qn = np.float32(np.linalg.norm(q))
    n = len(embeddings)
    best_s = np.empty(0, dtype=np.float32)
    best_r = np.empty(0, dtype=np.int64)
    for lo in range(0, n, block_rows):
        hi = min(n, lo + block_rows)
        blk = np.asarray(embeddings[lo:hi], dtype=np.float32)
        sims = blk @ q
        ok = np.ones(hi - lo, dtype=bool) if keep is None else np.asarray(keep[lo:hi], dtype=bool).copy()
        if mode == MODE_COSINE:
            norms = _row_norms(blk)
            ok &= norms > 0
            with np.errstate(divide="ignore", invalid="ignore"):
                sims = sims / (qn * norms)
        idx = np.flatnonzero(ok)
        if idx.size == 0:
            continue
        s = sims[idx].astype(np.float32)
        if idx.size > k:
            part = np.argpartition(-s, k - 1)[:k]
            # keep every row tied with the k-th so the boundary rule below decides
            kth = s[part].min()
            part = np.flatnonzero(s >= kth)
            s, idx = s[part], idx[part]
        best_s = np.concatenate([best_s, s])
        best_r = np.concatenate([best_r, idx.astype(np.int64) + lo + row_offset])
        if best_s.size > 4 * k:
            order = np.lexsort((-best_r, -best_s))[:k]
            best_s, best_r = best_s[order], best_r[order]
    order = np.lexsort((-best_r, -best_s))[:k]
    return best_s[order], best_r[order]


def query_blocks(
    queries: np.ndarray,
    blocks,
    k: int,
    *,
    mode: str = MODE_COSINE,
) -> List[Tuple[np.ndarray, np.ndarray]]:
    """:func:`query_vectorised` for several queries over a corpus that arrives block by block (the 10M-100M row
    configurations do not fit in host memory as fp32): ``blocks`` yields ``(first_row, rows[, keep])`` with rows
    float16 or float32.  Same arithmetic per row -- fp32 dot, fp32 norms, ``dot / (|q| * |e|)`` -- with the row
    norms computed once per block and the dots of all queries taken in one ``rows @ Q.T``; same ordering rule
    (similarity descending, then row descending).  Returns one ``(similarities, rows)`` per query."""
    Q = np.atleast_2d(np.asarray(queries, dtype=np.float32))
    qn = np.linalg.norm(Q, axis=1).astype(np.float32)
    best = [(np.empty(0, np.float32), np.empty(0, np.int64)) for _ in range(len(Q))]
    for item in blocks:
        first, blk = int(item[0]), np.asarray(item[1], dtype=np.float32)
        keep = None if len(item) < 3 or item[2] is None else np.asarray(item[2], dtype=bool)
        dots = blk @ Q.T
        ok = np.ones(len(blk), dtype=bool) if keep is None else keep.copy()
        norms = None
        if mode == MODE_COSINE:
            norms = _row_norms(blk)
            ok &= norms > 0
        idx = np.flatnonzero(ok)
        if idx.size == 0:
            continue
        for b in range(len(Q)):
            s = dots[idx, b]
            if mode == MODE_COSINE:
                s = s / (qn[b] * norms[idx])
            s = s.astype(np.float32)
            ids = idx
            if ids.size > k:
                part = np.argpartition(-s, k - 1)[:k]
                sel = np.flatnonzero(s >= s[part].min())       # every row tied with the k-th stays in
                s, ids = s[sel], ids[sel]
            bs = np.concatenate([best[b][0], s])
            br = np.concatenate([best[b][1], ids.astype(np.int64) + first])
            order = np.lexsort((-br, -bs))[:k]
            best[b] = (bs[order], br[order])
    return best


def score_rows(query_embedding: np.ndarray, rows: np.ndarray, mode: str = MODE_COSINE) -> np.ndarray:
    """The literal per-row fp32 formula applied to the given rows (used by the tie-aware
    comparator to look up the oracle score of a row the CUDA path returned)."""
    q = np.asarray(query_embedding, dtype=np.float32)
    out = np.empty(len(rows), dtype=np.float32)
    for i, e in enumerate(np.asarray(rows, dtype=np.float32)):
        out[i] = similarity(q, e, mode)
    return out


def score_rows_f64(query_embedding: np.ndarray, rows: np.ndarray, mode: str = MODE_COSINE) -> np.ndarray:
    """fp64 scores, only to classify near-ties and bound the fp32 oracle's own error."""
    q = np.asarray(query_embedding, dtype=np.float64)
    e = np.asarray(rows, dtype=np.float64)
    dots = e @ q
    if mode == MODE_DOT:
        return dots
    return dots / (np.linalg.norm(q) * np.linalg.norm(e, axis=1))


def lc_score(cosine: float) -> float:
    """LangChain ``Cassandra`` / cassio score convention for metric "cos": ``0.5 + 0.5*cos``
    (restated; call sites ``ingest/src/app/services/cassandra_service.py:79-83``)."""
    return 0.5 + 0.5 * float(cosine)

"""CPU oracle for the CodeRAG retrieval hot path.  TEST INFRASTRUCTURE ONLY.

This package is the *checker*, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it.  Nothing under ``githubrepostorag_b200/`` imports it, and the
product path raises if its CUDA library is missing rather than falling back to this.

PARITY UNPINNED.  The reference tree (jasonbuchanan145/GithubReposToRag) contains no
similarity arithmetic of its own and no retrieval test, golden vector or fixture
(SURVEY.md §0.1, §0.4, §8c).  The arithmetic lives in un-vendored third-party
packages with floor-only versions (``llama-index-core>=0.10.0``,
``langchain-community>=0.2.10``, ``cassio>=0.1.6``, ``graph-retriever>=0.2.0``;
``environment-worker.yaml:23-29,38-40``) and a Cassandra 5.0.3 server
(``helm/values.yaml:9-13``); none is importable or installable here.  The oracle
therefore restates the *published* algorithm of the library the north star names
(LlamaIndex in-memory similarity top-k in numpy fp32) and is anchored on the
reference's call sites (``rest_api/src/app/health.py:104-107``,
``rag_worker/src/worker/services/graph_rag_retrievers.py:68-80,104-134``).  The
golden fixtures in ``tests/golden/`` were produced by this restatement
(``tests/golden/make_golden.py``) plus hand-derived expected answers.
"""
from .simple_vector_store import (  # noqa: F401
    MODE_COSINE,
    MODE_DOT,
    similarity,
    prefilter_mask,
    filter_passes,
    get_top_k_embeddings,
    query_literal,
    query_vectorised,
    query_blocks,
    score_rows,
    score_rows_f64,
    lc_score,
)

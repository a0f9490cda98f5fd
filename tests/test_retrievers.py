"""The worker-facing factory (mirror of graph_rag_retrievers.py) and the Eager traversal.  The CPU
tests drive the traversal over an oracle-backed stand-in store; the GPU test runs the same traversal
over the real store and must return the same documents."""
import numpy as np
import pytest

import oracle
from githubrepostorag_b200 import synthetic as S
from githubrepostorag_b200.types import Content


class OracleStore:
    """Stand-in with the Adapter surface, answering from the oracle (test infrastructure)."""

    def __init__(self, emb16, metas, embedder):
        self.emb16, self.metas, self.embeddings = emb16, metas, embedder
        self.calls = []

    def _content(self, r):
        return Content(id=S.row_id(r), content=f"text {r}", embedding=self.emb16[r].astype(np.float32).tolist(),
                       metadata=dict(self.metas[r]))

    def _keep(self, flt):
        return oracle.prefilter_mask(self.metas, flt)

    def search_with_embedding(self, query, k=4, filter=None, **kw):
        q = np.asarray(self.embeddings.embed_query(query), np.float32)
        s, r = oracle.query_vectorised(q, self.emb16, k, keep=self._keep(filter))
        return q.tolist(), [self._content(int(i)) for i in r]

    def adjacent(self, edges, query_embedding, k, filter=None, **kw):
        self.calls.append(sorted(edges))
        q = np.asarray(query_embedding, np.float32)
        keep = np.zeros(len(self.metas), bool)
        for fld, val in edges:
            keep |= self._keep({**(filter or {}), fld: val})
        s, r = oracle.query_vectorised(q, self.emb16, k, keep=keep)
        return [self._content(int(i)) for i in r]


def _corpus(n=600, d=32, seed=2):
    rng = np.random.default_rng(seed)
    emb16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    metas = [{"namespace": "default" if i % 5 else "acme", "repo": f"repo{i % 4}", "module": f"mod{i % 13}",
              "file_path": f"src/f{i % 57}.py", "scope": "chunk"} for i in range(n)]

    class Emb:
        def embed_query(self, text):
            return emb16[int(text)].astype(np.float32).tolist()

        def embed_documents(self, texts):
            return [self.embed_query(t) for t in texts]
    return emb16, metas, Emb()


def test_factory_mirrors_reference_signatures():
    from githubrepostorag_b200.retrievers import Eager, GraphRetrieverFactory, TableNames, make_graph_retriever_factory
    emb16, metas, emb = _corpus()
    fake = OracleStore(emb16, metas, emb)
    f = make_graph_retriever_factory(hosts=["cassandra"], keyspace="vector_store", username=None, password=None,
                                     tables=TableNames(repo="embeddings_repo", module="embeddings_module",
                                                       file="embeddings_file", chunk="embeddings"),
                                     embedding_model="intfloat/e5-small-v2", store_loader=lambda t: fake)
    assert isinstance(f, GraphRetrieverFactory)
    r = f.for_chunk(k=10, start_k=3, adjacent_k=8, max_depth=2)          # agent_graph.py:175
    assert r.strategy == Eager(k=10, start_k=3, adjacent_k=8, max_depth=2)
    assert r.edges == [("namespace", "namespace"), ("repo", "repo"), ("module", "module"), ("file_path", "file_path")]
    assert f.for_repo(k=10, start_k=2, max_depth=2).edges == [("namespace", "namespace"), ("repo", "repo")]
    assert f.for_module().strategy.adjacent_k == 6 and f.for_file().strategy.k == 8
    assert f._store("embeddings") is fake and f._store("embeddings") is fake          # cached per table


def test_eager_traversal_over_oracle_store():
    from githubrepostorag_b200.retrievers import Eager, EagerGraphRetriever
    emb16, metas, emb = _corpus()
    st = OracleStore(emb16, metas, emb)
    r = EagerGraphRetriever(st, [("repo", "repo"), ("module", "module")], Eager(k=10, start_k=3, adjacent_k=8, max_depth=2))
    docs = r.invoke("17", filter={"namespace": "default"})
    assert len(docs) == 10 and docs[0].id == S.row_id(17) and len({d.id for d in docs}) == 10
    assert [d.metadata["_depth"] for d in docs[:3]] == [0, 0, 0] and all(d.metadata["_depth"] >= 1 for d in docs[3:])
    assert all(d.metadata["namespace"] == "default" for d in docs)
    assert all(isinstance(d.metadata["score"], float) and d.metadata["score"] == d.metadata["_similarity_score"] for d in docs)
    assert abs(docs[0].metadata["score"] - 1.0) < 1e-6
    start_meta = [metas[int(d.id[4:])] for d in docs[:3]]
    want_edges = sorted({("repo", m["repo"]) for m in start_meta} | {("module", m["module"]) for m in start_meta})
    assert st.calls[0] == want_edges                                   # one batched adjacent() per depth
    # depth-1 rows share a repo or module with a start row
    for d in docs[3:]:
        if d.metadata["_depth"] == 1:
            assert any(d.metadata["repo"] == m["repo"] or d.metadata["module"] == m["module"] for m in start_meta)
    assert EagerGraphRetriever(st, [], Eager(k=5, start_k=2)).invoke("3") and len(EagerGraphRetriever(st, [], Eager(k=5, start_k=2)).invoke("3")) == 2
    assert r.invoke("17", filter={"namespace": "nope"}) == []


@pytest.mark.gpu
def test_traversal_over_b200_store_equals_oracle_store():
    from githubrepostorag_b200.retrievers import GraphRetrieverFactory, TableNames
    from githubrepostorag_b200.vector_store import B200VectorStore
    emb16, metas, emb = _corpus(n=3000, d=64, seed=9)
    real = B200VectorStore(emb, dim=64)
    real.add_embeddings([S.row_id(i) for i in range(len(emb16))], emb16.astype(np.float32),
                        [f"text {i}" for i in range(len(emb16))], metas)
    fake = OracleStore(emb16, metas, emb)
    tables = TableNames(repo="r", module="m", file="f", chunk="c")
    f_real = GraphRetrieverFactory(tables=tables, embedding=emb, store_loader=lambda t: real)
    f_fake = GraphRetrieverFactory(tables=tables, embedding=emb, store_loader=lambda t: fake)
    for build in ("for_repo", "for_module", "for_file", "for_chunk"):
        for q, flt in (("5", {"namespace": "default"}), ("1234", None), ("77", {"namespace": "acme", "repo": "repo0"})):
            a = getattr(f_real, build)().invoke(q, filter=flt)
            b = getattr(f_fake, build)().invoke(q, filter=flt)
            assert [d.id for d in a] == [d.id for d in b], (build, q)
            assert [d.metadata["_depth"] for d in a] == [d.metadata["_depth"] for d in b]
            assert np.allclose([d.metadata["score"] for d in a], [d.metadata["score"] for d in b], atol=1e-6)
            assert all(d.page_content == f"text {int(d.id[4:])}" for d in a)
    f_real.close()

"""The retriever factory the worker builds its four scope retrievers from, over B200 stores.

Mirror of ``rag_worker/src/worker/services/graph_rag_retrievers.py`` (same names, arguments and
defaults): ``TableNames``, ``GraphRetrieverFactory(...).for_repo / for_module / for_file / for_chunk``,
``close()``, ``make_graph_retriever_factory(...)``.  ``GraphAgent`` (``agent_graph.py:151-176``) calls
exactly these and then ``retriever.invoke(q, filter=filters)`` (``:249,268``).

The reference returns ``langchain_graph_retriever.GraphRetriever(store, edges, strategy=Eager(...))``.
With that package installed the same class can be built over a :class:`B200VectorStore` (it accepts
any graph-retriever ``Adapter``; the store implements ``search_with_embedding / search / get /
adjacent``).  It is not installed here, so :class:`EagerGraphRetriever` restates the traversal of
``graph_retriever.traverse`` with the ``Eager`` strategy -- UPSTREAM (recalled, graph-retriever >= 0.2.0,
``environment-worker.yaml:27-28``):

1. ``start_k`` nearest rows under the caller's filter (depth 0);
2. every newly discovered row is selected (Eager), until ``k`` rows are selected;
3. the outgoing edges of the newly selected rows -- ``(target_field, row.metadata[source_field])`` for
   each configured edge -- that were not followed before are followed: ``adjacent(edges, qvec,
   k=adjacent_k, filter)``, which here is ONE batched device search for all edges;
4. rows not seen before enter at depth + 1; repeat while depth < ``max_depth`` and rows are missing.

Documents carry ``_depth`` and ``_similarity_score`` (what graph-retriever writes) plus ``score``
(what the worker's ``_score_of`` reads, ``agent_graph.py:44-61``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from .types import Document
from .vector_store import B200VectorStore


@dataclass(frozen=True)
class TableNames:
    repo: str
    module: str
    file: str
    chunk: str


@dataclass(frozen=True)
class Eager:
    """graph-retriever's ``Eager`` strategy parameters (defaults as upstream)."""
    k: int = 5
    start_k: int = 4
    adjacent_k: int = 10
    max_depth: Optional[int] = None


class EagerGraphRetriever:
    """``GraphRetriever(store=..., edges=..., strategy=Eager(...))`` over a B200 store."""

    def __init__(self, store: B200VectorStore, edges: Sequence[Tuple[str, str]], strategy: Eager):
        self.store = store
        self.edges = [(str(a), str(b)) for a, b in edges]
        self.strategy = strategy

    @staticmethod
    def _cos(q: np.ndarray, e: Sequence[float]) -> float:
        v = np.asarray(e, dtype=np.float32)
        den = float(np.linalg.norm(q) * np.linalg.norm(v))
        return float(np.dot(q, v) / den) if den > 0 else 0.0

    def invoke(self, query: str, filter: Optional[Dict[str, Any]] = None, **kwargs: Any) -> List[Document]:
        st = self.strategy
        k = int(kwargs.get("k", st.k))
        qvec, found = self.store.search_with_embedding(query, k=st.start_k, filter=filter, as_arrays=True)
        q = np.asarray(qvec, dtype=np.float32)
        seen: Dict[str, Tuple[Any, int]] = {}
        selected: List[str] = []
        followed = set()
        frontier = []
        for c in found:
            if c.id not in seen:
                seen[c.id] = (c, 0)
                frontier.append(c)
        depth = 0
        while frontier and len(selected) < k:
            take = frontier[: k - len(selected)]
            selected.extend(c.id for c in take)
            if len(selected) >= k or (st.max_depth is not None and depth >= st.max_depth):
                break
            new_edges = []
            for c in take:
                for src, dst in self.edges:
                    val = c.metadata.get(src)
                    if val is None:
                        continue
                    for v in (val if isinstance(val, (list, tuple, set)) else [val]):
                        e = (dst, str(v))
                        if e not in followed:
                            followed.add(e)
                            new_edges.append(e)
            if not new_edges:
                break
            depth += 1
            frontier = []
            for c in self.store.adjacent(new_edges, q, k=st.adjacent_k, filter=filter, as_arrays=True):
                if c.id not in seen:
                    seen[c.id] = (c, depth)
                    frontier.append(c)
        docs = []
        # one matrix product for all selected rows (graph-retriever scores each node from its embedding)
        sims = {}
        if selected:
            E = np.stack([np.asarray(seen[rid][0].embedding, dtype=np.float32) for rid in selected])
            den = np.linalg.norm(q) * np.linalg.norm(E, axis=1)
            dots = E @ q
            sims = {rid: (float(dots[i] / den[i]) if den[i] > 0 else 0.0) for i, rid in enumerate(selected)}
        for rid in selected:
            c, d = seen[rid]
            sim = sims[rid]
            md = dict(c.metadata)
            md.update({"_depth": d, "_similarity_score": sim, "score": sim})
            docs.append(Document(page_content=c.content, metadata=md, id=c.id))
        return docs

    get_relevant_documents = invoke


class GraphRetrieverFactory:
    """Builds one retriever per scope over per-table B200 stores sharing one embedder.

    ``store_loader(table_name) -> B200VectorStore`` supplies the resident table (e.g. migrated with
    ``migrate.load_cassandra_rows`` or restored with ``B200VectorStore.load``); ``hosts / keyspace /
    username / password`` are accepted for signature compatibility with the reference factory.
    """

    def __init__(self, *, tables: TableNames, embedding: Any = None, embedding_model: str = "intfloat/e5-small-v2",
                 store_loader: Optional[Callable[[str], B200VectorStore]] = None, device: int = 0, dim: int = 384,
                 hosts: Optional[List[str]] = None, keyspace: Optional[str] = None, username: Optional[str] = None,
                 password: Optional[str] = None) -> None:
        self.hosts, self.keyspace, self.username, self.password = hosts, keyspace, username, password
        self.tables = tables
        self.embedding_model = embedding_model
        self._emb = embedding
        self._loader = store_loader
        self._device, self._dim = device, dim
        self._stores: Dict[str, B200VectorStore] = {}

    def _scope_of(self, table_name: str) -> Optional[str]:
        """The writer's scope of one of this factory's tables (``vector_write_service.py:37-43``): it decides
        which metadata keys the table is filterable on."""
        for scope, name in (("repo", self.tables.repo), ("module", self.tables.module), ("file", self.tables.file),
                            ("chunk", self.tables.chunk)):
            if name == table_name:
                return scope
        return None

    def _store(self, table_name: str) -> B200VectorStore:
        st = self._stores.get(table_name)
        if st:
            return st
        if self._loader is not None:
            st = self._loader(table_name)
            if getattr(st, "embeddings", None) is None:
                st.embeddings = self._emb
        else:
            st = B200VectorStore(self._emb, dim=self._dim, table_name=table_name, device=self._device,
                                 scope=self._scope_of(table_name))
        self._stores[table_name] = st
        return st

    @staticmethod
    def _edges_for(scope: str):
        if scope == "project":
            return [("namespace", "namespace"), ("repo", "repo")]
        if scope == "package":
            return [("namespace", "namespace"), ("repo", "repo"), ("module", "module")]
        return [("namespace", "namespace"), ("repo", "repo"), ("module", "module"), ("file_path", "file_path")]

    def for_repo(self, *, k: int = 6, start_k: int = 2, max_depth: int = 2):
        return EagerGraphRetriever(self._store(self.tables.repo), self._edges_for("project"),
                                   Eager(k=k, start_k=start_k, max_depth=max_depth))

    def for_module(self, *, k: int = 8, start_k: int = 2, adjacent_k: int = 6, max_depth: int = 2):
        return EagerGraphRetriever(self._store(self.tables.module), self._edges_for("package"),
                                   Eager(k=k, start_k=start_k, adjacent_k=adjacent_k, max_depth=max_depth))

    def for_file(self, *, k: int = 8, start_k: int = 2, adjacent_k: int = 6, max_depth: int = 2):
        return EagerGraphRetriever(self._store(self.tables.file), self._edges_for("file"),
                                   Eager(k=k, start_k=start_k, adjacent_k=adjacent_k, max_depth=max_depth))

    def for_chunk(self, *, k: int = 10, start_k: int = 3, adjacent_k: int = 8, max_depth: int = 2):
        return EagerGraphRetriever(self._store(self.tables.chunk), self._edges_for("code"),
                                   Eager(k=k, start_k=start_k, adjacent_k=adjacent_k, max_depth=max_depth))

    def close(self):
        for st in self._stores.values():
            st.close()
        self._stores.clear()


def make_graph_retriever_factory(*, hosts: Optional[List[str]] = None, keyspace: Optional[str] = None,
                                 username: Optional[str] = None, password: Optional[str] = None,
                                 tables: TableNames, embedding_model: str = "intfloat/e5-small-v2",
                                 **b200_kwargs: Any) -> GraphRetrieverFactory:
    return GraphRetrieverFactory(hosts=hosts, keyspace=keyspace, username=username, password=password,
                                 tables=tables, embedding_model=embedding_model, **b200_kwargs)

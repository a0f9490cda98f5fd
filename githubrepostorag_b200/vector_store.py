"""``B200VectorStore`` -- the drop-in object behind the reference's retrievers.

The reference's boundary for the retrieval hot path is a duck-typed Python store handed to the
retriever (``rag_worker/src/worker/services/graph_rag_retrievers.py:68-80,104-134``).  This
class exports the three faces SURVEY.md §8b lists, all served by the same CUDA path:

* **A -- LlamaIndex ``VectorStore``**: ``add / delete / query(VectorStoreQuery) ->
  VectorStoreQueryResult`` and ``as_retriever(similarity_top_k=k).retrieve(q)``
  (``rest_api/src/app/health.py:104-107``).  Similarities are raw cosine, descending.
* **B -- LangChain ``Cassandra`` store**: ``similarity_search_with_embedding_id_by_vector``,
  ``similarity_search_with_score_id_by_vector`` (score ``0.5 + 0.5*cos``), ``similarity_search``,
  ``similarity_search_with_score``, ``add_documents``, ``get_by_document_id``, ``as_retriever``
  (``ingest/src/app/services/cassandra_service.py:79-86``; ``vector_write_service.py:159``).
* **B' -- graph-retriever ``Adapter``**: ``search_with_embedding / search / get / adjacent``;
  ``adjacent`` runs the whole fan-out (one query vector, one filter per edge) as ONE search call.

Error behaviour follows the reference: failures raise (callers catch broadly,
``agent_graph.py:276-278``); fewer than k survivors returns the short list.  Host-side ids,
text and metadata stay in Python, indexed by the row numbers the device returns.
"""
from __future__ import annotations

import threading
from typing import Any, Callable, Dict, Iterable, List, Mapping, Optional, Sequence, Tuple

import numpy as np

from . import _native as N
from .index import DeviceShard, MetadataDictionary
from .types import (Content, Document, FilterCondition, FilterOperator, MetadataFilter, MetadataFilters, NodeWithScore,
                    TextNode, VectorStoreQuery, VectorStoreQueryResult)

# keys the reference filters or traverses on (agent_graph.py:200-233; graph_rag_retrievers.py:82-100;
# north star adds branch / component_kind, written at ingest_controller.py:169-173)
DEFAULT_FILTER_KEYS = ("namespace", "repo", "module", "file_path", "branch", "component_kind", "scope", "language")


class _Rows:
    """Host-side id / text / metadata of the rows, list-backed or synthesised on demand."""

    def __init__(self):
        self.ids: List[str] = []
        self.texts: List[str] = []
        self.metas: List[Dict[str, Any]] = []
        self.pos: Dict[str, int] = {}
        self.id_fn: Optional[Callable[[int], str]] = None
        self.text_fn: Optional[Callable[[int], str]] = None
        self.meta_fn: Optional[Callable[[int], Dict[str, Any]]] = None
        self.n_lazy = 0

    def __len__(self) -> int:
        return self.n_lazy if self.id_fn else len(self.ids)

    def id(self, r: int) -> str:
        return self.id_fn(r) if self.id_fn else self.ids[r]

    def text(self, r: int) -> str:
        if self.id_fn:
            return self.text_fn(r) if self.text_fn else ""
        return self.texts[r]

    def meta(self, r: int) -> Dict[str, Any]:
        if self.id_fn:
            return dict(self.meta_fn(r)) if self.meta_fn else {}
        return dict(self.metas[r])


class B200VectorStore:
    stores_text: bool = True
    is_embedding_query: bool = True
    flat_metadata: bool = True

    def __init__(self, embedding: Any = None, *, dim: Optional[int] = None,
                 filter_keys: Sequence[str] = DEFAULT_FILTER_KEYS, device: int = 0,
                 table_name: str = "embeddings", metric: str = "cosine"):
        N.load()  # fail loudly at construction when the CUDA library is missing
        self.embeddings = embedding          # object with embed_query / embed_documents
        self.table_name = table_name
        self.metric = metric
        self.device = device
        self.dim = dim
        self.dictionary = MetadataDictionary(filter_keys)
        self._rows = _Rows()
        self._pending: List[np.ndarray] = []
        self._codes: List[List[int]] = []
        self._shard: Optional[DeviceShard] = None
        self._dirty = False
        self._lock = threading.RLock()

    # ------------------------------------------------------------------ construction / writes
    @property
    def client(self) -> "B200VectorStore":
        return self

    @classmethod
    def from_arrays(cls, embeddings: np.ndarray, codes: Optional[np.ndarray], dictionary: MetadataDictionary,
                    *, id_fn: Callable[[int], str], meta_fn: Optional[Callable[[int], Dict[str, Any]]] = None,
                    text_fn: Optional[Callable[[int], str]] = None, embedding: Any = None, device: int = 0,
                    metric: str = "cosine", row_offset: int = 0) -> "B200VectorStore":
        """Bulk load (synthetic corpora, migrations): rows + pre-coded metadata columns; ids,
        text and metadata are produced on demand from the row number."""
        st = cls(embedding, dim=int(embeddings.shape[1]), filter_keys=dictionary.keys, device=device, metric=metric)
        st.dictionary = dictionary
        st._rows.id_fn, st._rows.meta_fn, st._rows.text_fn = id_fn, meta_fn, text_fn
        st._rows.n_lazy = int(embeddings.shape[0])
        st._shard = DeviceShard.from_host(embeddings, codes, device=device, row_offset=row_offset)
        return st

    def _append(self, ids: Sequence[str], embeddings: np.ndarray, texts: Sequence[str],
                metadatas: Sequence[Mapping[str, Any]]) -> List[str]:
        emb = np.asarray(embeddings, dtype=np.float32)
        if emb.ndim != 2 or len(ids) != emb.shape[0]:
            raise ValueError("embeddings must be [len(ids), dim]")
        with self._lock:
            if self._rows.id_fn:
                raise RuntimeError("this store was bulk-loaded with from_arrays and is read-only")
            if self.dim is None:
                self.dim = int(emb.shape[1])
            if emb.shape[1] != self.dim:
                raise ValueError(f"embedding dimension {emb.shape[1]} != store dimension {self.dim}")
            fresh_rows = []
            for i, rid in enumerate(ids):
                rid = str(rid)
                md = {str(k): v for k, v in dict(metadatas[i] or {}).items() if v is not None}
                codes = self.dictionary.encode_row(md)
                if rid in self._rows.pos:      # upsert, as a Cassandra INSERT on the primary key
                    r = self._rows.pos[rid]
                    self._rows.texts[r], self._rows.metas[r] = texts[i], md
                    self._codes[r] = codes
                    self._overwrite(r, emb[i])
                else:
                    self._rows.pos[rid] = len(self._rows.ids)
                    self._rows.ids.append(rid)
                    self._rows.texts.append(texts[i])
                    self._rows.metas.append(md)
                    self._codes.append(codes)
                    fresh_rows.append(i)
            if fresh_rows:
                self._pending.append(emb[fresh_rows].astype(np.float16))
            self._dirty = True
        return [str(i) for i in ids]

    def _overwrite(self, r: int, vec: np.ndarray) -> None:
        lo = 0
        for blk in self._pending:
            if r < lo + len(blk):
                blk[r - lo] = vec.astype(np.float16)
                return
            lo += len(blk)

    def add(self, nodes: Sequence[Any], **kwargs: Any) -> List[str]:
        """LlamaIndex face: nodes carry ``embedding``, ``node_id``, ``metadata``, text."""
        ids, embs, texts, metas = [], [], [], []
        for n in nodes:
            if getattr(n, "embedding", None) is None:
                raise ValueError("node without embedding")
            ids.append(getattr(n, "node_id", None) or getattr(n, "id_", None))
            embs.append(n.embedding)
            try:
                texts.append(n.get_content(metadata_mode="none"))
            except Exception:
                texts.append(getattr(n, "text", "") or "")
            metas.append(dict(getattr(n, "metadata", {}) or {}))
        return self._append(ids, np.asarray(embs, dtype=np.float32), texts, metas) if ids else []

    def add_embeddings(self, ids: Sequence[str], embeddings: np.ndarray, texts: Optional[Sequence[str]] = None,
                       metadatas: Optional[Sequence[Mapping[str, Any]]] = None) -> List[str]:
        n = len(ids)
        return self._append(ids, embeddings, texts or [""] * n, metadatas or [{}] * n)

    def add_texts(self, texts: Iterable[str], metadatas: Optional[List[dict]] = None,
                  ids: Optional[List[str]] = None, **kwargs: Any) -> List[str]:
        texts = list(texts)
        if self.embeddings is None:
            raise RuntimeError("add_texts needs an embedder (embedding=...)")
        embs = np.asarray(self.embeddings.embed_documents(texts), dtype=np.float32)
        if ids is None or len(ids) == 0:
            import uuid
            ids = [uuid.uuid4().hex for _ in texts]
        return self._append(ids, embs, texts, metadatas or [{}] * len(texts))

    def add_documents(self, documents: Sequence[Any], ids: Optional[Sequence[str]] = None, **kwargs: Any) -> List[str]:
        """LangChain face used by the ingest writer (``vector_write_service.py:158-159``)."""
        docs = list(documents)
        return self.add_texts([d.page_content for d in docs], [dict(d.metadata or {}) for d in docs],
                              list(ids) if ids else None)

    def delete(self, ref_doc_id: str, **kwargs: Any) -> None:
        with self._lock:
            r = self._rows.pos.get(str(ref_doc_id))
            if r is None:
                return
            emb = np.concatenate(self._pending) if self._pending else np.empty((0, self.dim or 0), np.float16)
            keep = np.ones(len(emb), dtype=bool)
            keep[r] = False
            self._pending = [emb[keep]]
            for lst in (self._rows.ids, self._rows.texts, self._rows.metas, self._codes):
                del lst[r]
            self._rows.pos = {rid: i for i, rid in enumerate(self._rows.ids)}
            self._dirty = True

    def _ensure_built(self) -> DeviceShard:
        with self._lock:
            if self._shard is not None and not self._dirty:
                return self._shard
            if self._rows.id_fn:
                return self._shard  # bulk-loaded, never dirty
            if self.dim is None:
                raise RuntimeError("empty store: dimension unknown (pass dim=...)")
            emb = np.concatenate(self._pending) if self._pending else np.empty((0, self.dim), np.float16)
            self._pending = [emb]
            codes = (np.asarray(self._codes, dtype=np.int32).T.copy() if self._codes
                     else np.empty((self.dictionary.n_keys, 0), np.int32))
            if self._shard is not None:
                self._shard.close()
            self._shard = DeviceShard.from_host(emb, codes, device=self.device)
            self._dirty = False
            return self._shard

    def close(self) -> None:
        """Release device resources (mirrors ``GraphRetrieverFactory.close``, ``graph_rag_retrievers.py:136-140``)."""
        with self._lock:
            if self._shard is not None:
                self._shard.close()
                self._shard = None

    def __len__(self) -> int:
        return len(self._rows)

    # ------------------------------------------------------------------ the hot path
    def _search(self, queries: np.ndarray, k: int, filters: Sequence[Optional[Mapping[str, Any]]],
                algo: int = 0) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        shard = self._ensure_built()
        k = int(k)
        if k < 1:
            raise ValueError("k must be >= 1")
        if k > N.MAX_TOPK:
            raise ValueError(f"k={k} exceeds the supported maximum {N.MAX_TOPK}")
        filt = None
        if any(f for f in filters):
            filt = np.stack([self.dictionary.encode_filter(f) for f in filters])
        return shard.search_host(queries, k, metric=self.metric, filt=filt, algo=algo)

    def _embed_query(self, text: str) -> List[float]:
        if self.embeddings is None:
            raise RuntimeError("text query needs an embedder (embedding=...)")
        return list(self.embeddings.embed_query(text))

    @staticmethod
    def _filters_to_dict(filters: Any) -> Optional[Dict[str, Any]]:
        if filters is None:
            return None
        if isinstance(filters, Mapping):
            return dict(filters)
        cond = getattr(filters, "condition", FilterCondition.AND)
        if cond is not None and str(getattr(cond, "value", cond)).lower() != "and":
            raise NotImplementedError("only AND of equality filters is supported (what the reference emits)")
        out: Dict[str, Any] = {}
        for f in getattr(filters, "filters", []) or []:
            op = getattr(f, "operator", FilterOperator.EQ)
            if str(getattr(op, "value", op)) != "==":
                raise NotImplementedError(f"filter operator {op!r} is not supported (EQ only)")
            if f.key in out and out[f.key] != f.value:
                return {"__impossible__": "1"}  # k == a AND k == b
            out[f.key] = f.value
        return out

    # ------------------------------------------------------------------ face A: LlamaIndex
    def query(self, query: VectorStoreQuery, **kwargs: Any) -> VectorStoreQueryResult:
        mode = getattr(query, "mode", None)
        if mode is not None and str(getattr(mode, "value", mode)) != "default":
            raise NotImplementedError(f"query mode {mode!r} is not supported (DEFAULT only)")
        if getattr(query, "node_ids", None):
            raise NotImplementedError("node_ids restriction is not supported")
        if query.query_embedding is None:
            raise ValueError("query_embedding is required (is_embedding_query=True)")
        q = np.asarray(query.query_embedding, dtype=np.float32)[None, :]
        flt = self._filters_to_dict(getattr(query, "filters", None))
        scores, rows, n_valid = self._search(q, query.similarity_top_k, [flt])
        nodes, sims, ids = [], [], []
        for j in range(int(n_valid[0])):
            r = int(rows[0, j])
            rid = self._rows.id(r)
            nodes.append(TextNode(text=self._rows.text(r), metadata=self._rows.meta(r), id_=rid))
            sims.append(float(scores[0, j]))
            ids.append(rid)
        return VectorStoreQueryResult(nodes=nodes, similarities=sims, ids=ids)

    async def aquery(self, query: VectorStoreQuery, **kwargs: Any) -> VectorStoreQueryResult:
        return self.query(query, **kwargs)

    # ------------------------------------------------------------------ face B: LangChain Cassandra
    def _doc(self, r: int, cos: float) -> Document:
        md = self._rows.meta(r)
        # the worker reads `_score`/`score` first (agent_graph.py:44-61); graph-retriever writes
        # `_similarity_score` -- populate both so sources carry a numeric score
        # (rest_api/src/app/static/index.html:97 calls score.toFixed)
        md["score"] = float(cos)
        md["_similarity_score"] = float(cos)
        return Document(page_content=self._rows.text(r), metadata=md, id=self._rows.id(r))

    def _hits(self, embedding: Sequence[float], k: int, filter: Optional[Mapping[str, Any]]):
        q = np.asarray(embedding, dtype=np.float32)[None, :]
        scores, rows, n_valid = self._search(q, k, [dict(filter) if filter else None])
        return [(int(rows[0, j]), float(scores[0, j])) for j in range(int(n_valid[0]))]

    def similarity_search_with_embedding_id_by_vector(
            self, embedding: List[float], k: int = 4, filter: Optional[Dict[str, str]] = None,
            body_search: Any = None) -> List[Tuple[Document, List[float], str]]:
        if body_search:
            raise NotImplementedError("body_search is not supported")
        hits = self._hits(embedding, k, filter)
        vecs = self._ensure_built().fetch_rows([r for r, _ in hits])
        return [(self._doc(r, s), vecs[i].tolist(), self._rows.id(r)) for i, (r, s) in enumerate(hits)]

    def similarity_search_with_score_id_by_vector(
            self, embedding: List[float], k: int = 4, filter: Optional[Dict[str, str]] = None,
            body_search: Any = None) -> List[Tuple[Document, float, str]]:
        if body_search:
            raise NotImplementedError("body_search is not supported")
        return [(self._doc(r, s), 0.5 + 0.5 * s, self._rows.id(r)) for r, s in self._hits(embedding, k, filter)]

    def similarity_search_with_score_by_vector(self, embedding, k: int = 4, filter=None, body_search=None):
        return [(d, s) for d, s, _ in self.similarity_search_with_score_id_by_vector(embedding, k, filter, body_search)]

    def similarity_search_by_vector(self, embedding, k: int = 4, filter=None, body_search=None, **kw):
        return [d for d, _, _ in self.similarity_search_with_score_id_by_vector(embedding, k, filter, body_search)]

    def similarity_search(self, query: str, k: int = 4, filter: Optional[Dict[str, str]] = None, **kw) -> List[Document]:
        return self.similarity_search_by_vector(self._embed_query(query), k, filter)

    def similarity_search_with_score(self, query: str, k: int = 4, filter: Optional[Dict[str, str]] = None, **kw):
        return self.similarity_search_with_score_by_vector(self._embed_query(query), k, filter)

    def get_by_document_id(self, document_id: str) -> Optional[Document]:
        r = self._row_of(document_id)
        return None if r is None else Document(page_content=self._rows.text(r), metadata=self._rows.meta(r), id=str(document_id))

    def _row_of(self, rid: str) -> Optional[int]:
        if self._rows.id_fn:
            try:
                r = int(str(rid).rsplit("-", 1)[-1])
            except ValueError:
                return None
            return r if 0 <= r < len(self._rows) and self._rows.id(r) == str(rid) else None
        return self._rows.pos.get(str(rid))

    def as_retriever(self, **kwargs: Any) -> "B200Retriever":
        return B200Retriever(self, **kwargs)

    # ------------------------------------------------------------------ face B': graph-retriever Adapter
    def _content(self, r: int, vec: np.ndarray) -> Content:
        return Content(id=self._rows.id(r), content=self._rows.text(r), embedding=vec.tolist(),
                       metadata=self._rows.meta(r))

    def search_with_embedding(self, query: str, k: int = 4, filter: Optional[Dict[str, Any]] = None, **kw
                              ) -> Tuple[List[float], List[Content]]:
        qvec = self._embed_query(query)
        return qvec, self.search(qvec, k=k, filter=filter)

    def search(self, embedding: Sequence[float], k: int = 4, filter: Optional[Dict[str, Any]] = None, **kw) -> List[Content]:
        hits = self._hits(embedding, k, filter)
        vecs = self._ensure_built().fetch_rows([r for r, _ in hits])
        return [self._content(r, vecs[i]) for i, (r, _) in enumerate(hits)]

    def get(self, ids: Sequence[str], filter: Optional[Dict[str, Any]] = None, **kw) -> List[Content]:
        rows = []
        for rid in dict.fromkeys(ids):
            r = self._row_of(rid)
            if r is None:
                continue
            if filter and any(str(self._rows.meta(r).get(k)) != str(v) for k, v in filter.items()):
                continue
            rows.append(r)
        vecs = self._ensure_built().fetch_rows(rows)
        return [self._content(r, vecs[i]) for i, r in enumerate(rows)]

    def adjacent(self, edges: Iterable[Any], query_embedding: Sequence[float], k: int,
                 filter: Optional[Dict[str, Any]] = None, **kw) -> List[Content]:
        """All rows adjacent through ``edges`` -- one filtered top-k per distinct
        ``(field, value)`` edge, same query vector -- issued as ONE batched search (the
        reference's adapter issues one CQL query per edge), then merged:
        de-duplicated by row, best ``k`` by similarity."""
        pairs = []
        for e in edges:
            if isinstance(e, tuple):
                fld, val = e[0], e[1]
            else:
                fld, val = getattr(e, "incoming_field", getattr(e, "field", None)), getattr(e, "value", None)
            if fld is None:
                raise ValueError(f"unsupported edge {e!r}")
            pairs.append((str(fld), val))
        pairs = list(dict.fromkeys(pairs))
        if not pairs:
            return []
        base = dict(filter or {})
        flts = []
        for fld, val in pairs:
            if fld in base and str(base[fld]) != str(val):
                flts.append({"__impossible__": "1"})
            else:
                flts.append({**base, fld: val})
        q = np.repeat(np.asarray(query_embedding, dtype=np.float32)[None, :], len(flts), axis=0)
        scores, rows, n_valid = self._search(q, k, flts)
        best: Dict[int, float] = {}
        for i in range(len(flts)):
            for j in range(int(n_valid[i])):
                r = int(rows[i, j])
                s = float(scores[i, j])
                if r not in best or s > best[r]:
                    best[r] = s
        top = sorted(best.items(), key=lambda t: (-t[1], -t[0]))[:k]
        vecs = self._ensure_built().fetch_rows([r for r, _ in top])
        return [self._content(r, vecs[i]) for i, (r, _) in enumerate(top)]


class B200Retriever:
    """``store.as_retriever(...)``: LangChain-style ``invoke(query, filter=...)`` and
    LlamaIndex-style ``retrieve(query)`` (``health.py:106-107``)."""

    def __init__(self, store: B200VectorStore, *, similarity_top_k: Optional[int] = None,
                 search_kwargs: Optional[Dict[str, Any]] = None, filters: Any = None, **_: Any):
        self.store = store
        sk = dict(search_kwargs or {})
        self.k = int(similarity_top_k or sk.get("k", 4))
        self.filter = sk.get("filter") or B200VectorStore._filters_to_dict(filters)

    def invoke(self, query: str, filter: Optional[Dict[str, Any]] = None, **kw) -> List[Document]:
        flt = {**(self.filter or {}), **(filter or {})} or None
        return self.store.similarity_search(query, k=self.k, filter=flt)

    get_relevant_documents = invoke

    def retrieve(self, query: Any) -> List[NodeWithScore]:
        text = getattr(query, "query_str", query)
        emb = getattr(query, "embedding", None) or self.store._embed_query(text)
        res = self.store.query(VectorStoreQuery(query_embedding=emb, similarity_top_k=self.k,
                                                filters=MetadataFilters(filters=[MetadataFilter(key=k, value=v) for k, v in self.filter.items()])
                                                if self.filter else None))
        return [NodeWithScore(node=n, score=s) for n, s in zip(res.nodes or [], res.similarities or [])]

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

# list-valued metadata the ingest writer shreds into key->value entries for traversal
# (ingest/src/app/services/vector_write_service.py:28-34,152-154): a filter on one of these keys
# means MEMBERSHIP ("the row's list contains the value"), served by a per-query row mask.
DEFAULT_MULTI_VALUED_KEYS = ("topics", "imports", "labels")
# The writer's per-scope allow-lists of INDEXED (= filterable) metadata keys, as in the reference
# (ingest/src/app/services/vector_write_service.py:28-34) ...
ALLOW_FIELDS_BY_SCOPE = {
    "catalog": ("namespace", "repo", "owner", "language", "topics", "labels", "component_kind"),
    "repo": ("namespace", "repo", "owner", "language", "topics", "labels"),
    "module": ("namespace", "repo", "module", "language", "topics", "imports", "labels"),
    "file": ("namespace", "repo", "module", "file_path", "language", "topics", "imports", "labels"),
    "chunk": ("namespace", "repo", "module", "file_path", "symbol", "language", "topics", "imports"),
}
# ... and the keys it keeps on every row whatever the scope (vector_write_service.py:58-64; `row_id` is the
# primary key, not a filter).  Stored keys are made filterable too: a superset of what SAI indexes, so every
# filter the reference can run (agent_graph.py:200-233; graph_rag_retrievers.py:82-100; the north star's
# branch / component_kind, ingest_controller.py:169-173) runs here.
KEEP_ALWAYS_KEYS = ("scope", "namespace", "repo", "module", "file_path", "symbol", "owner", "component_kind",
                    "branch", "language")
# scope <-> table (vector_write_service.py:37-43; agent_graph.py:163-168)
TABLES_BY_SCOPE = {"catalog": "embeddings_catalog", "repo": "embeddings_repo", "module": "embeddings_module",
                   "file": "embeddings_file", "chunk": "embeddings"}


def filter_keys_for_scope(scope: Optional[str]) -> Tuple[Tuple[str, ...], Tuple[str, ...]]:
    """(single-valued coded keys, list-valued keys) a table of ``scope`` filters on: the scope's allow-list
    plus the always-kept keys; ``None`` / unknown scope = the union over all scopes."""
    allow: List[str] = []
    for sc in ([scope] if scope in ALLOW_FIELDS_BY_SCOPE else list(ALLOW_FIELDS_BY_SCOPE)):
        allow += [k for k in ALLOW_FIELDS_BY_SCOPE[sc] if k not in allow]
    coded = [k for k in allow if k not in DEFAULT_MULTI_VALUED_KEYS]
    coded += [k for k in KEEP_ALWAYS_KEYS if k not in coded]
    multi = [k for k in allow if k in DEFAULT_MULTI_VALUED_KEYS]
    return tuple(coded), tuple(multi)


def scope_of_table(table_name: str) -> Optional[str]:
    for sc, tb in TABLES_BY_SCOPE.items():
        if tb == table_name:
            return sc
    return None


# every key the reference can filter or traverse on, whatever the table
DEFAULT_FILTER_KEYS = filter_keys_for_scope(None)[0]


def _retrieval_histogram():
    """``rag_b200_retrieval_seconds``: the search call's latency, next to the worker's own
    ``rag_worker_retrieval_seconds`` (``rag_worker/src/worker/worker.py:47``).  None when
    prometheus_client is not installed."""
    try:
        from prometheus_client import REGISTRY, Histogram
        existing = getattr(REGISTRY, "_names_to_collectors", {}).get("rag_b200_retrieval_seconds")
        if existing is not None:
            return existing
        return Histogram("rag_b200_retrieval_seconds", "B200 vector search latency (one store call, all its queries)",
                         ["table", "kernel_path"],
                         buckets=(1e-4, 2.5e-4, 5e-4, 1e-3, 2.5e-3, 5e-3, 1e-2, 2.5e-2, 5e-2, 0.1, 0.25, 1.0))
    except Exception:  # noqa: BLE001
        return None


class _WriteGate:
    """Many concurrent searches (the worker's thread pool) or one writer (append / delete /
    compact re-bind the device buffers, which no search may be using).  Writers have preference:
    once one is waiting, new readers queue behind it (re-entrant reads of a thread that is already
    reading pass), so a stream of searches cannot starve an ingest."""

    def __init__(self):
        self._cv = threading.Condition()
        self._readers = 0
        self._writer: Optional[int] = None
        self._depth = 0
        self._waiting_writers = 0
        self._local = threading.local()
        self._read_hold = _ReadHold(self)
        self._write_hold = _WriteHold(self)

    def read(self) -> "_ReadHold":
        return self._read_hold

    def write(self) -> "_WriteHold":
        return self._write_hold

    def _enter_read(self) -> None:
        me = threading.get_ident()
        depth = getattr(self._local, "reads", 0)
        with self._cv:
            if depth == 0 and self._writer != me:
                while self._writer is not None or self._waiting_writers > 0:
                    self._cv.wait()
            self._readers += 1
        self._local.reads = depth + 1

    def _exit_read(self) -> None:
        self._local.reads -= 1
        with self._cv:
            self._readers -= 1
            if self._readers == 0 and self._waiting_writers:
                self._cv.notify_all()

    def _enter_write(self) -> None:
        me = threading.get_ident()
        with self._cv:
            if self._writer == me:
                self._depth += 1
                return
            if getattr(self._local, "reads", 0):
                raise RuntimeError("internal: write requested while this thread holds a read")
            self._waiting_writers += 1
            try:
                while self._writer is not None or self._readers > 0:
                    self._cv.wait()
            finally:
                self._waiting_writers -= 1
            self._writer, self._depth = me, 1

    def _exit_write(self) -> None:
        with self._cv:
            self._depth -= 1
            if self._depth == 0:
                self._writer = None
                self._cv.notify_all()


class _ReadHold:
    """``with gate.read():`` -- stateless (the gate keeps the counts), so one object serves every thread."""
    __slots__ = ("_gate",)

    def __init__(self, gate: _WriteGate):
        self._gate = gate

    def __enter__(self):
        self._gate._enter_read()

    def __exit__(self, *exc):
        self._gate._exit_read()


class _WriteHold:
    __slots__ = ("_gate",)

    def __init__(self, gate: _WriteGate):
        self._gate = gate

    def __enter__(self):
        self._gate._enter_write()

    def __exit__(self, *exc):
        self._gate._exit_write()


class _Rows:
    """Host-side id / text / metadata of the rows, list-backed or synthesised on demand."""

    def __init__(self):
        self.ids: List[Optional[str]] = []      # None = deleted row (tombstoned on the device)
        self.texts: List[str] = []
        self.metas: List[Dict[str, Any]] = []
        self.pos: Dict[str, int] = {}
        self.n_dead = 0
        self.id_fn: Optional[Callable[[int], str]] = None
        self.text_fn: Optional[Callable[[int], str]] = None
        self.meta_fn: Optional[Callable[[int], Dict[str, Any]]] = None
        self.n_lazy = 0

    def __len__(self) -> int:
        return self.n_lazy if self.id_fn else len(self.ids) - self.n_dead

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


def _reading(fn):
    """Run a read-side face method under the store's read gate, after staged writes are flushed, so
    the row numbers a search returns stay valid while ids / text / vectors are looked up."""
    import functools

    @functools.wraps(fn)
    def wrapped(self, *a, **kw):
        self._ensure_built()
        with self._gate.read():
            return fn(self, *a, **kw)
    return wrapped


class B200VectorStore:
    stores_text: bool = True
    is_embedding_query: bool = True
    flat_metadata: bool = True

    def __init__(self, embedding: Any = None, *, dim: Optional[int] = None,
                 filter_keys: Optional[Sequence[str]] = None, device: int = 0,
                 table_name: str = "embeddings", metric: str = "cosine",
                 multi_valued_keys: Optional[Sequence[str]] = None, scope: Optional[str] = None):
        N.load()  # fail loudly at construction when the CUDA library is missing
        # filterable keys: given explicitly, else the writer's lists for the table's scope (``scope=...``; the
        # retriever factory and the migration loader pass it), else the union over all scopes
        sc_coded, sc_multi = filter_keys_for_scope(scope)
        filter_keys = sc_coded if filter_keys is None else filter_keys
        multi_valued_keys = (sc_multi if scope else DEFAULT_MULTI_VALUED_KEYS) if multi_valued_keys is None else multi_valued_keys
        self.scope = scope
        self.embeddings = embedding          # object with embed_query / embed_documents
        self.table_name = table_name
        self.metric = metric
        self.device = device
        self.dim = dim
        self.dictionary = MetadataDictionary(filter_keys)
        self._rows = _Rows()
        self._shard: Optional[DeviceShard] = None
        self._gate = _WriteGate()
        # writes are staged on the host and reach the device (append / tombstone) at the next read
        self._pending_emb: List[np.ndarray] = []
        self._pending_codes: List[np.ndarray] = []
        self._pending_dead: List[int] = []
        # membership postings of list-valued metadata: (key, value) -> host row numbers
        self.multi_valued_keys = tuple(k for k in multi_valued_keys if k not in self.dictionary.keys)
        self._postings: Dict[Tuple[str, str], List[int]] = {}
        self._mask_cache: Dict[Any, Any] = {}
        self._filter_cache: Dict[Any, Any] = {}
        self._mask_lock = threading.Lock()
        self._hist = _retrieval_histogram()

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

    @classmethod
    def from_shard(cls, shard: DeviceShard, dictionary: Optional[MetadataDictionary] = None, *,
                   id_fn: Callable[[int], str], meta_fn: Optional[Callable[[int], Dict[str, Any]]] = None,
                   text_fn: Optional[Callable[[int], str]] = None, embedding: Any = None, metric: str = "cosine"
                   ) -> "B200VectorStore":
        """A read-only store over rows that are already resident (a shard built on the device); ids, text and
        metadata are produced on demand from the LOCAL row number."""
        st = cls(embedding, dim=shard.dim, filter_keys=dictionary.keys if dictionary is not None else (),
                 device=shard.device.index or 0, metric=metric)
        if dictionary is not None:
            st.dictionary = dictionary
        st._rows.id_fn, st._rows.meta_fn, st._rows.text_fn = id_fn, meta_fn, text_fn
        st._rows.n_lazy = shard.n_rows
        st._shard = shard
        return st

    def _append(self, ids: Sequence[str], embeddings: np.ndarray, texts: Sequence[str],
                metadatas: Sequence[Mapping[str, Any]]) -> List[str]:
        """Upsert rows.  New rows are encoded straight into the device shard's spare capacity
        (``DeviceShard.append`` -- no re-upload of what is already resident); a row whose id
        exists is tombstoned and its replacement appended, as a CQL INSERT on the primary key
        overwrites."""
        emb = np.asarray(embeddings, dtype=np.float32)
        if emb.ndim != 2 or len(ids) != emb.shape[0]:
            raise ValueError("embeddings must be [len(ids), dim]")
        with self._gate.write():
            if self._rows.id_fn:
                raise RuntimeError("this store was bulk-loaded with from_arrays and is read-only")
            if self.dim is None:
                self.dim = int(emb.shape[1])
            if emb.shape[1] != self.dim:
                raise ValueError(f"embedding dimension {emb.shape[1]} != store dimension {self.dim}")
            last = {str(rid): i for i, rid in enumerate(ids)}          # duplicates inside the batch: last wins
            order = sorted(last.values())
            dead, codes = [], []
            base = len(self._rows.ids)
            for j, i in enumerate(order):
                rid = str(ids[i])
                md = {str(k): v for k, v in dict(metadatas[i] or {}).items() if v is not None}
                old = self._rows.pos.get(rid)
                if old is not None:
                    dead.append(old)
                    self._rows.ids[old] = None
                    self._rows.n_dead += 1
                self._rows.pos[rid] = base + j
                self._rows.ids.append(rid)
                self._rows.texts.append(texts[i])
                self._rows.metas.append(md)
                codes.append(self.dictionary.encode_row(md))
                self._post(base + j, md)
            code_mat = np.asarray(codes, dtype=np.int32).T.copy() if codes else np.empty((self.dictionary.n_keys, 0), np.int32)
            self._pending_emb.append(emb[order].astype(np.float16))
            self._pending_codes.append(code_mat)
            self._pending_dead.extend(dead)
        return [str(i) for i in ids]

    @staticmethod
    def _members(v: Any) -> List[str]:
        if isinstance(v, (list, tuple, set, frozenset)):
            return [str(x) for x in v]
        return [str(v)]

    def _post(self, row: int, md: Mapping[str, Any]) -> None:
        for key in self.multi_valued_keys:
            if key in md:
                for val in self._members(md[key]):
                    self._postings.setdefault((key, val), []).append(row)

    def _flush(self) -> None:
        """Bring staged writes to the device: append the new rows, tombstone the replaced / deleted."""
        with self._gate.write():
            self._mask_cache.clear()
            if self._pending_emb:
                new_emb = np.concatenate(self._pending_emb)
                code_mat = np.concatenate(self._pending_codes, axis=1)
                self._pending_emb, self._pending_codes = [], []
                if self._shard is None:
                    self._shard = DeviceShard.from_host(new_emb, code_mat, device=self.device,
                                                        capacity=max(1024, 2 * len(new_emb)))
                else:
                    self._shard.append(new_emb, code_mat)
            if self._pending_dead and self._shard is not None:
                dead = [r for r in self._pending_dead if r < self._shard.n_rows]
                self._pending_dead = []
                if dead:
                    self._shard.tombstone(dead)
            self._maybe_compact()

    def _maybe_compact(self) -> None:
        """Rebuild the shard without its tombstones once they outnumber the live rows."""
        r = self._rows
        if r.n_dead > 1024 and r.n_dead > len(r.ids) - r.n_dead:
            self.compact()

    def compact(self, cluster_by: Optional[Sequence[str]] = None) -> None:
        """Drop tombstoned rows (device-side gather of the live rows into fresh buffers).

        ``cluster_by``: also ORDER the rows by these coded keys, outermost first (``("repo", "module", "file_path")``;
        ties keep their insertion order).  Rows that share a code then form one contiguous run, the shard's per-code
        row ranges (taken when the columns are bound) become exact, and a search filtered on such keys -- every
        traversal edge of the reference is one -- is an unfiltered streaming pass over the run: no filter pass, no
        gather.  Row numbers change, ids do not; rows appended later land at the end and merely widen a range."""
        import torch
        with self._gate.write():
            r = self._rows
            if self._pending_emb or self._pending_dead:
                self._flush()
            if r.id_fn or self._shard is None or (r.n_dead == 0 and not cluster_by):
                return
            live = [i for i, rid in enumerate(r.ids) if rid is not None]
            old = self._shard
            if cluster_by and live and old.codes is not None:
                cols = []
                for key in cluster_by:
                    pos = self.dictionary._pos.get(key)
                    if pos is None:
                        raise ValueError(f"cluster_by: {key!r} is not a coded key of this store ({self.dictionary.keys})")
                    cols.append(old.codes[pos][:old.n_rows].cpu().numpy()[live])
                order = np.lexsort(tuple(reversed(cols)))            # last key of lexsort is the primary one
                live = [live[i] for i in order]
            idx = torch.as_tensor(live, dtype=torch.int64, device=self._shard.device)
            rows = old.rows.index_select(0, idx).contiguous()
            inv = old.inv_norm.index_select(0, idx).contiguous()
            codes = None if old.codes is None else [c.index_select(0, idx).contiguous() for c in old.codes]
            if rows.shape[0] == 0:
                rows = torch.empty((0, old.dim), dtype=torch.float16, device=old.device)
            self._shard = DeviceShard(rows, inv, codes, old.row_offset)
            old.close()
            r.ids = [r.ids[i] for i in live]
            r.texts = [r.texts[i] for i in live]
            r.metas = [r.metas[i] for i in live]
            r.pos = {rid: i for i, rid in enumerate(r.ids)}
            r.n_dead = 0
            self._postings = {}
            self._mask_cache.clear()
            for i, md in enumerate(r.metas):
                self._post(i, md)

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
        """Tombstone the row on the device (``crb_tombstone_rows``); unknown ids are ignored."""
        with self._gate.write():
            r = self._rows.pos.pop(str(ref_doc_id), None)
            if r is None:
                return
            self._rows.ids[r] = None
            self._rows.n_dead += 1
            self._pending_dead.append(r)

    def _ensure_built(self) -> DeviceShard:
        if self._pending_emb or self._pending_dead:
            self._flush()
        if self._shard is None:
            if self.dim is None:
                raise RuntimeError("empty store: dimension unknown (pass dim=...)")
            with self._gate.write():
                if self._shard is None:
                    self._shard = DeviceShard.from_host(np.empty((0, self.dim), np.float16),
                                                        np.empty((self.dictionary.n_keys, 0), np.int32),
                                                        device=self.device, capacity=1024)
        return self._shard

    def close(self) -> None:
        """Release device resources (mirrors ``GraphRetrieverFactory.close``, ``graph_rag_retrievers.py:136-140``)."""
        with self._gate.write():
            if self._shard is not None:
                self._shard.close()
                self._shard = None

    # ------------------------------------------------------------------ persistence
    SAVE_BLOCK_ROWS = 1 << 18

    def save(self, path: str) -> None:
        """Checkpoint the store (live rows only).  The reference's equivalent is the Cassandra PVC
        (``helm/values.yaml:25-28``).  A path ending in ``.npz`` writes one file (small tables: everything passes
        through host memory at once); any other path is a DIRECTORY written in a streaming fashion -- the fp16 rows
        and code columns go from the device to ``rows16.npy`` / ``codes.npy`` in 2**18-row blocks through a memory map,
        ids / text / metadata as one JSON line per row -- so a 10M-row table needs a block of host memory, not the
        table."""
        import json
        import os
        with self._gate.write():
            if self._rows.id_fn:
                raise RuntimeError("bulk-loaded (from_arrays) stores keep no ids/text to save")
            self.compact() if self._rows.n_dead else None
            sh = self._ensure_built()
            head = {"format": 2, "dim": self.dim, "metric": self.metric, "table_name": self.table_name, "scope": self.scope,
                    "keys": self.dictionary.keys, "values": self.dictionary.values, "n_rows": sh.n_rows,
                    "multi_valued_keys": list(self.multi_valued_keys)}
            if str(path).endswith(".npz"):
                meta = dict(head, format=1, ids=self._rows.ids, texts=self._rows.texts, metas=self._rows.metas)
                blob = np.frombuffer(json.dumps(meta).encode("utf-8"), dtype=np.uint8)
                codes = np.stack([c.cpu().numpy().astype(np.int32) for c in sh.codes]) if sh.codes is not None and sh.n_keys \
                    else np.empty((0, sh.n_rows), np.int32)
                with open(path, "wb") as f:
                    np.savez(f, rows16=sh.rows.cpu().numpy(), codes=codes, meta=blob)
                return
            os.makedirs(path, exist_ok=True)
            n, blk = sh.n_rows, self.SAVE_BLOCK_ROWS
            rows_mm = np.lib.format.open_memmap(os.path.join(path, "rows16.npy"), mode="w+", dtype=np.float16, shape=(n, self.dim))
            codes_mm = np.lib.format.open_memmap(os.path.join(path, "codes.npy"), mode="w+", dtype=np.int32, shape=(sh.n_keys, n))
            for lo in range(0, n, blk):
                hi = min(n, lo + blk)
                rows_mm[lo:hi] = sh.rows[lo:hi].cpu().numpy()
                for m in range(sh.n_keys):
                    codes_mm[m, lo:hi] = sh.codes[m][lo:hi].cpu().numpy()
            rows_mm.flush(); codes_mm.flush()
            del rows_mm, codes_mm
            with open(os.path.join(path, "rows.jsonl"), "w", encoding="utf-8") as f:
                for rid, text, md in zip(self._rows.ids, self._rows.texts, self._rows.metas):
                    f.write(json.dumps({"id": rid, "text": text, "metadata": md}) + "\n")
            with open(os.path.join(path, "meta.json"), "w", encoding="utf-8") as f:
                json.dump(head, f)

    @staticmethod
    def read_saved(path: str):
        """``(head, rows16, codes, ids, texts, metas)`` of a saved table; the two arrays are memory-mapped for the
        directory format, so a caller can take a row range without reading the rest (the sharded loader does)."""
        import json
        import os
        if os.path.isdir(path):
            with open(os.path.join(path, "meta.json"), encoding="utf-8") as f:
                head = json.load(f)
            rows16 = np.load(os.path.join(path, "rows16.npy"), mmap_mode="r")
            codes = np.load(os.path.join(path, "codes.npy"), mmap_mode="r")
            ids, texts, metas = [], [], []
            with open(os.path.join(path, "rows.jsonl"), encoding="utf-8") as f:
                for line in f:
                    rec = json.loads(line)
                    ids.append(rec["id"]); texts.append(rec["text"]); metas.append(rec["metadata"])
            return head, rows16, codes, ids, texts, metas
        with np.load(path) as z:
            head = json.loads(bytes(z["meta"]).decode("utf-8"))
            rows16, codes = z["rows16"], z["codes"]
        return head, rows16, codes, list(head.pop("ids")), list(head.pop("texts")), list(head.pop("metas"))

    @classmethod
    def load(cls, path: str, *, embedding: Any = None, device: int = 0) -> "B200VectorStore":
        head, rows16, codes, ids, texts, metas = cls.read_saved(path)
        st = cls(embedding, dim=int(head["dim"]), filter_keys=head["keys"], device=device,
                 table_name=head.get("table_name", "embeddings"), metric=head.get("metric", "cosine"),
                 multi_valued_keys=head.get("multi_valued_keys"), scope=head.get("scope"))
        st.dictionary.values = [dict(v) for v in head["values"]]
        st._rows.ids, st._rows.texts, st._rows.metas = ids, texts, metas
        st._rows.pos = {rid: i for i, rid in enumerate(st._rows.ids)}
        for i, md in enumerate(st._rows.metas):
            st._post(i, md)
        # (from_host uploads in 2**18-row blocks, so a memory-mapped table never sits in host memory whole)
        st._shard = DeviceShard.from_host(rows16, np.asarray(codes) if codes.shape[0] else None, device=device,
                                          capacity=max(1024, len(rows16)))
        return st

    def __len__(self) -> int:
        return len(self._rows)

    # ------------------------------------------------------------------ the hot path
    def _search(self, queries: np.ndarray, k: int, filters: Sequence[Optional[Mapping[str, Any]]],
                algo: int = 0, vectors: bool = False):
        """``(scores, rows, n_valid)`` -- and, with ``vectors``, the hits' stored vectors fp32 ``[B, k, dim]`` as a fourth
        result, brought back by the search's own device-to-host copy where the plan allows."""
        k = int(k)
        if k < 1:
            raise ValueError("k must be >= 1")
        if k > N.MAX_TOPK:
            raise ValueError(f"k={k} exceeds the supported maximum {N.MAX_TOPK}")
        compiled = [self._compile_filter(f) for f in filters]
        split = [(eq, extra) for eq, extra, _ in compiled]
        filt = None
        if any(row is not None for _, _, row in compiled):
            filt = np.full((len(compiled), self.dictionary.n_keys), N.FILT_ANY, dtype=np.int32)
            for i, (_, _, row) in enumerate(compiled):
                if row is not None:
                    filt[i] = row
        import time
        t0 = time.perf_counter()
        try:
            res = self._search_locked(queries, k, filt, split, algo, vectors)
            if vectors and len(res) == 3:       # this plan could not carry them: one fetch for all hits
                s_, r_, v_ = res
                vecs = np.zeros((r_.shape[0], k, self.dim), dtype=np.float32)
                where = [(b, j) for b in range(r_.shape[0]) for j in range(int(v_[b]))]
                if where:
                    got = self._fetch_vectors([int(r_[b, j]) for b, j in where])
                    for (b, j), vec in zip(where, got):
                        vecs[b, j] = vec
                res = (s_, r_, v_, vecs)
            return res
        finally:
            if self._hist is not None:
                self._hist.labels(self.table_name, "masked" if any(e for _, e in split) else "host").observe(time.perf_counter() - t0)

    def _search_locked(self, queries, k, filt, split, algo, vectors=False):
        with self._gate.read():
            vecs = None
            if not any(extra for _, extra in split):
                got = self._shard.search_host(queries, k, metric=self.metric, filt=filt, algo=algo, vectors=vectors)
                s_, r_, v_ = got[:3]
                vecs = got[3] if vectors else None
            else:
                # predicates one code per key cannot express ride on a per-query row mask (device path)
                import torch
                masks = [self._row_mask(extra) if extra else None for _, extra in split]
                q = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32)).to(self._shard.device)
                s_, r_, v_ = self._shard.search_device(q, k, metric=self.metric, filt=filt, algo=algo, row_masks=masks)
                s_, r_, v_ = s_.cpu().numpy(), r_.cpu().numpy(), v_.cpu().numpy()
            # the device returns GLOBAL rows (row_offset + local); everything host-side is indexed by LOCAL row
            off = self._shard.row_offset
            if off:
                r_ = np.where(r_ >= 0, r_ - off, r_)
            return (s_, r_, v_) if vecs is None else (s_, r_, v_, vecs)

    # ---- filter vocabulary --------------------------------------------------------------------------
    # A filter is a flat dict, ANDed (the only shape the reference emits, agent_graph.py:198-239):
    #   key: value                 equality on a coded key; MEMBERSHIP on a list-valued key
    #   key: ("in", [values])      IN on a coded key; "shares a member with" on a list-valued key
    #   "__any_of__": [group, ..]  each group a list of alternative filter dicts: OR inside a group, AND across
    #                              groups (LlamaIndex FilterCondition.OR and nested MetadataFilters)
    @staticmethod
    def _canon(flt: Mapping[str, Any]) -> tuple:
        """Hashable canonical form: sorted ``(key, ("eq", v) | ("in", (v, ...)))`` items; ``__any_of__`` groups
        as tuples of canonical alternatives."""
        items = []
        for key, val in flt.items():
            if key == "__any_of__":
                items.append((key, tuple(tuple(B200VectorStore._canon(alt) for alt in group) for group in val)))
            elif isinstance(val, tuple) and len(val) == 2 and val[0] == "in":
                items.append((str(key), ("in", tuple(sorted(str(v) for v in val[1])))))
            else:
                items.append((str(key), ("eq", str(val))))
        return tuple(sorted(items, key=repr))

    def _compile_filter(self, flt: Optional[Mapping[str, Any]]):
        """filter dict -> ``(eq, extra, codes)``: :meth:`_split_filter`'s two parts and the equalities as one int32
        code per key (None without equalities).  Remembered per distinct filter: the worker repeats a handful of
        filters per request (``agent_graph.py:198-239``) and ``adjacent`` one per edge.  A filter naming a value the
        dictionary does not know yet is not remembered (a later write may introduce the value)."""
        if not flt:
            return None, (), None
        try:
            key = tuple(flt.items())
            hit = self._filter_cache.get(key)
        except TypeError:                      # an unhashable value (("in", [..]) lists, __any_of__ groups)
            key, hit = None, None
        if hit is not None:
            return hit
        eq, extra = self._split_filter(flt)
        row = self.dictionary.encode_filter(eq) if eq else None
        hit = (eq, extra, row)
        if key is not None and (row is None or int(row.min()) > N.FILT_IMPOSSIBLE):
            if len(self._filter_cache) >= 4096:
                self._filter_cache.clear()
            self._filter_cache[key] = hit
        return hit

    def _split_filter(self, flt: Optional[Mapping[str, Any]]):
        """filter dict -> (equalities on coded keys -- they go to the filter kernel as codes --, extra
        predicates in canonical form -- they ride on a row mask)."""
        if not flt:
            return None, ()
        eq: Dict[str, Any] = {}
        extra = []
        for key, spec in self._canon(flt):
            if key != "__any_of__" and spec[0] == "eq" and key not in self.multi_valued_keys:
                eq[key] = spec[1]
            else:
                extra.append((key, spec))
        return (eq or None), tuple(extra)

    def _keep_items(self, items) -> "torch.Tensor":
        """bool [n_rows] on the device: rows satisfying every canonical predicate of ``items``."""
        import torch
        sh = self._shard
        keep = torch.ones(sh.n_rows, dtype=torch.bool, device=sh.device)
        for key, spec in items:
            if key == "__any_of__":
                for group in spec:
                    union = torch.zeros(sh.n_rows, dtype=torch.bool, device=sh.device)
                    for alt in group:
                        union |= self._keep_items(alt)
                    keep &= union
                continue
            values = [spec[1]] if spec[0] == "eq" else list(spec[1])
            if key in self.multi_valued_keys:
                # membership in the row's list: the union of the values' postings
                rows = sorted({r for v in values for r in self._postings.get((key, v), ()) if r < sh.n_rows})
                hit = torch.zeros(sh.n_rows, dtype=torch.bool, device=sh.device)
                if rows:
                    hit[torch.as_tensor(rows, dtype=torch.int64, device=sh.device)] = True
                keep &= hit
                continue
            pos = self.dictionary._pos.get(key)
            codes = [] if pos is None else [self.dictionary.values[pos][v] for v in values if v in self.dictionary.values[pos]]
            if not codes or sh.codes is None:
                keep.zero_()                         # unknown key or value: no row has it
            elif len(codes) == 1:
                keep &= sh.codes[pos] == codes[0]
            else:
                keep &= torch.isin(sh.codes[pos], torch.as_tensor(codes, dtype=sh.codes[pos].dtype, device=sh.device))
        return keep

    def _row_mask(self, extra):
        """Device pass-bit mask of the rows satisfying every extra predicate (cached until the next write)."""
        with self._mask_lock:
            cached = self._mask_cache.get(extra)
        if cached is not None:
            return cached
        mask = self._shard.pack_row_mask(self._keep_items(extra))
        with self._mask_lock:
            if len(self._mask_cache) > 256:
                self._mask_cache.clear()
            self._mask_cache[extra] = mask
        return mask

    def _fetch_vectors(self, rows: Sequence[int]) -> np.ndarray:
        """Stored vectors of host rows, upcast to fp32 (the ``embedding`` of LCCassandra's result triples)."""
        return self._shard.fetch_rows(rows)

    def search_rows(self, queries: np.ndarray, k: int, filters: Optional[Sequence[Optional[Mapping[str, Any]]]] = None
                    ) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        """The array-level call under all three faces: fp32 ``queries [B, dim]`` (host), one filter dict per query (or
        None) -> ``(scores fp32 [B, k], rows int64 [B, k], n_valid int32 [B])`` on the host; ``rows`` index this
        store's ids / text / metadata (``row_id(r)``).  One host-to-device copy of the queries and one device-to-host
        copy of the result inside."""
        self._ensure_built()
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        return self._search(q, k, list(filters) if filters is not None else [None] * len(q))

    def row_id(self, row: int) -> str:
        return self._rows.id(int(row))

    def _embed_query(self, text: str) -> List[float]:
        if self.embeddings is None:
            raise RuntimeError("text query needs an embedder (embedding=...)")
        return list(self.embeddings.embed_query(text))

    @staticmethod
    def _filters_to_dict(filters: Any) -> Optional[Dict[str, Any]]:
        """LlamaIndex ``MetadataFilters`` (EQ / IN / CONTAINS leaves, AND / OR conditions, nesting) -> filter dict."""
        if filters is None:
            return None
        if isinstance(filters, Mapping):
            return dict(filters)
        cond = getattr(filters, "condition", FilterCondition.AND)
        cond = "and" if cond is None else str(getattr(cond, "value", cond)).lower()
        if cond not in ("and", "or"):
            raise NotImplementedError(f"filter condition {cond!r} is not supported (AND, OR)")
        leaves, nested = [], []
        for f in getattr(filters, "filters", []) or []:
            if hasattr(f, "filters"):
                nested.append(B200VectorStore._filters_to_dict(f) or {})
                continue
            op = getattr(f, "operator", FilterOperator.EQ)
            name = str(getattr(op, "value", op)).lower()
            if name in ("==", "contains"):        # CONTAINS on a list-valued key == membership
                val: Any = f.value
            elif name == "in":
                val = ("in", list(f.value))
            else:
                raise NotImplementedError(f"filter operator {op!r} is not supported (EQ, IN, CONTAINS)")
            leaves.append((f.key, val))
        if cond == "or":
            alts = [{k: v} for k, v in leaves] + nested
            return {"__any_of__": [alts]} if alts else None
        out: Dict[str, Any] = {}
        for key, val in leaves:
            if key in out and out[key] != val:
                return {"__impossible__": "1"}  # k == a AND k == b
            out[key] = val
        if nested:
            out["__any_of__"] = [[nd] for nd in nested]
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
        flt = self._filters_to_dict(getattr(query, "filters", None))
        return self._query(np.asarray(query.query_embedding, dtype=np.float32)[None, :], int(query.similarity_top_k), flt)

    @_reading
    def _query(self, q: np.ndarray, k: int, flt: Optional[Dict[str, Any]]) -> VectorStoreQueryResult:
        scores, rows, n_valid = self._search(q, k, [flt])
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

    def _hits_with_vectors(self, embedding: Sequence[float], k: int, filter: Optional[Mapping[str, Any]]):
        """Hits and their stored vectors from ONE search call (the vectors ride on the result copy)."""
        q = np.asarray(embedding, dtype=np.float32)[None, :]
        scores, rows, n_valid, vecs = self._search(q, k, [dict(filter) if filter else None], vectors=True)
        nv = int(n_valid[0])
        return [(int(rows[0, j]), float(scores[0, j])) for j in range(nv)], vecs[0, :nv]

    @_reading
    def similarity_search_with_embedding_id_by_vector(
            self, embedding: List[float], k: int = 4, filter: Optional[Dict[str, str]] = None,
            body_search: Any = None) -> List[Tuple[Document, List[float], str]]:
        if body_search:
            raise NotImplementedError("body_search is not supported")
        hits, vecs = self._hits_with_vectors(embedding, k, filter)
        return [(self._doc(r, s), vecs[i].tolist(), self._rows.id(r)) for i, (r, s) in enumerate(hits)]

    @_reading
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
                g = int(str(rid).rsplit("-", 1)[-1])
            except ValueError:
                return None
            off = self._shard.row_offset if self._shard is not None else 0
            for r in (g, g - off):           # ids made from the local or from the global row number
                if 0 <= r < len(self._rows) and self._rows.id(r) == str(rid):
                    return r
            return None
        return self._rows.pos.get(str(rid))

    def as_retriever(self, **kwargs: Any) -> "B200Retriever":
        return B200Retriever(self, **kwargs)

    # ------------------------------------------------------------------ face B': graph-retriever Adapter
    def _content(self, r: int, vec: np.ndarray, as_arrays: bool = False) -> Content:
        return Content(id=self._rows.id(r), content=self._rows.text(r), embedding=vec if as_arrays else vec.tolist(),
                       metadata=self._rows.meta(r))

    def search_with_embedding(self, query: str, k: int = 4, filter: Optional[Dict[str, Any]] = None, **kw
                              ) -> Tuple[List[float], List[Content]]:
        qvec = self._embed_query(query)
        return qvec, self.search(qvec, k=k, filter=filter, **kw)

    @_reading
    def search(self, embedding: Sequence[float], k: int = 4, filter: Optional[Dict[str, Any]] = None, *,
               as_arrays: bool = False, **kw) -> List[Content]:
        """``as_arrays``: leave ``Content.embedding`` as the fp32 array the device returned instead of a list of
        floats (the in-repo traversal scores from arrays; graph-retriever's own callers get the list)."""
        hits, vecs = self._hits_with_vectors(embedding, k, filter)
        return [self._content(r, vecs[i], as_arrays) for i, (r, _) in enumerate(hits)]

    @_reading
    def get(self, ids: Sequence[str], filter: Optional[Dict[str, Any]] = None, **kw) -> List[Content]:
        rows = []
        for rid in dict.fromkeys(ids):
            r = self._row_of(rid)
            if r is None:
                continue
            if filter and any(str(self._rows.meta(r).get(k)) != str(v) for k, v in filter.items()):
                continue
            rows.append(r)
        vecs = self._fetch_vectors(rows)
        return [self._content(r, vecs[i]) for i, r in enumerate(rows)]

    MAX_ALTERNATIVES = 32            # distinct edge filters one crb_search_host_any_of call takes

    def _search_any_of(self, q: np.ndarray, k: int, flts: Sequence[Mapping[str, Any]]):
        """The edges of one ``adjacent`` step as ONE search over the union of their row sets: the best ``k`` of a union
        is the best ``k`` of the per-edge best ``k``'s (what the adapter computes), and a row in several sets is scored
        once.  ``(rows, vectors, n_valid)``, or None when an edge needs more than coded equalities (list-valued keys,
        IN, OR: those ride on row masks) -- the caller then issues the per-edge batch."""
        compiled = [self._compile_filter(f) for f in flts]
        if any(extra or row is None for _, extra, row in compiled):
            return None
        import time
        t0 = time.perf_counter()
        try:
            with self._gate.read():
                alts = np.concatenate([row for _, _, row in compiled]).reshape(len(compiled), -1)
                _, rows, nv, vecs = self._shard.search_host_any_of(q, alts, k, metric=self.metric, vectors=True)
                off = self._shard.row_offset
                if off:
                    rows = np.where(rows >= 0, rows - off, rows)
                return rows, vecs, nv
        finally:
            if self._hist is not None:
                self._hist.labels(self.table_name, "any_of").observe(time.perf_counter() - t0)

    @_reading
    def adjacent(self, edges: Iterable[Any], query_embedding: Sequence[float], k: int,
                 filter: Optional[Dict[str, Any]] = None, *, as_arrays: bool = False, **kw) -> List[Content]:
        """The best ``k`` rows adjacent through ``edges`` (the reference's adapter issues one CQL query per
        ``(field, value)`` edge with the same query vector, merges the hits, de-duplicates them and keeps the best
        ``k`` by similarity).  Edges that are coded equalities go out as ONE search over the union of
        their row sets (:meth:`_search_any_of` -- same rows, a row shared by several edges scored once); otherwise as
        one batched search, a query per distinct edge, merged here."""
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
        # edges that resolve to the same filter (an edge on a key the caller's filter already pins) are one search:
        # the results are merged by row below, and single-query groups keep the fan-out on the balanced kernel
        try:
            flts = list({tuple(f.items()): f for f in flts}.values())        # (base's items first, then the edge's: one order)
        except TypeError:                                                    # an unhashable value in the caller's filter
            flts = list({tuple(sorted((str(a), str(b)) for a, b in f.items())): f for f in flts}.values())
        qv = np.asarray(query_embedding, dtype=np.float32)
        union = self._search_any_of(qv, k, flts) if 2 <= len(flts) <= self.MAX_ALTERNATIVES else None
        if union is not None:
            rows1, vecs1, nv = union
            return [self._content(int(rows1[j]), vecs1[j], as_arrays) for j in range(nv)]
        q = np.repeat(qv[None, :], len(flts), axis=0)
        scores, rows, n_valid, vecs = self._search(q, k, flts, vectors=True)
        # merge the per-edge lists: one entry per row (its best score), best k by (score desc, row desc)
        valid = (np.arange(rows.shape[1])[None, :] < np.asarray(n_valid)[:, None]).ravel()
        slot = np.nonzero(valid)[0]
        if not len(slot):
            return []
        fr, fs = rows.ravel()[slot], scores.ravel()[slot]
        order = np.lexsort((-fs, fr))
        lead = np.ones(len(order), dtype=bool)
        lead[1:] = fr[order][1:] != fr[order][:-1]
        cand = order[lead]
        top = cand[np.lexsort((-fr[cand], -fs[cand]))[:k]]
        flat = vecs.reshape(-1, vecs.shape[-1])
        return [self._content(int(fr[t]), flat[slot[t]], as_arrays) for t in top]


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

"""Loader from the reference's Cassandra row format.

Every embeddings table has the shape ``(row_id TEXT PRIMARY KEY, attributes_blob TEXT, body_blob TEXT,
vector VECTOR<FLOAT,384>, metadata_s MAP<TEXT,TEXT>)`` (``helm/templates/cassandra-initdb-configmap.yaml:14-20,
86-92``); the worker reads one table per scope (``rag_worker/src/worker/services/agent_graph.py:163-168``).
``load_cassandra_rows`` feeds such rows -- dicts, named tuples or plain ``(row_id, body_blob, vector,
metadata_s)`` tuples, e.g. straight from ``session.execute("SELECT row_id, body_blob, vector, metadata_s FROM
vector_store.embeddings")`` -- into a :class:`~githubrepostorag_b200.vector_store.B200VectorStore`.
"""
from __future__ import annotations

from typing import Any, Iterable, Mapping

import numpy as np

from .vector_store import TABLES_BY_SCOPE, B200VectorStore, scope_of_table  # noqa: F401  (re-exported)


def store_for_table(table_name: str, *, dim: int = 384, embedding: Any = None, device: int = 0) -> B200VectorStore:
    """An empty store for one of the reference's embeddings tables, filterable on the keys the ingest writer
    indexes for that table's scope plus the keys it keeps on every row
    (``ingest/src/app/services/vector_write_service.py:28-43,58-64``)."""
    return B200VectorStore(embedding, dim=dim, table_name=table_name, device=device, scope=scope_of_table(table_name))


def _field(row: Any, name: str, pos: int):
    if isinstance(row, Mapping):
        return row.get(name)
    if hasattr(row, name):
        return getattr(row, name)
    return row[pos]


def load_cassandra_rows(store, rows: Iterable[Any], batch: int = 4096) -> int:
    """Upsert CQL rows into ``store`` in batches; returns the number of rows read.  Rows without a
    vector are skipped (the reference writes none, but a half-written row must not poison the index)."""
    ids, vecs, texts, metas, n = [], [], [], [], 0

    def flush():
        if ids:
            store.add_embeddings(list(ids), np.asarray(vecs, dtype=np.float32), list(texts), list(metas))
            ids.clear(); vecs.clear(); texts.clear(); metas.clear()

    for row in rows:
        n += 1
        vec = _field(row, "vector", 2)
        if vec is None:
            continue
        ids.append(str(_field(row, "row_id", 0)))
        texts.append(_field(row, "body_blob", 1) or "")
        vecs.append(np.asarray(vec, dtype=np.float32))
        metas.append(dict(_field(row, "metadata_s", 3) or {}))
        if len(ids) >= batch:
            flush()
    flush()
    return n

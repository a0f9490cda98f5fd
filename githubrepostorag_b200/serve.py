"""Multi-GPU serving entry point: one process per GPU, ONE store object.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \\
        -m githubrepostorag_b200.serve --load /data/embeddings.npz [--app worker.worker:serve]

Every rank loads its row range of the table saved by ``B200VectorStore.save`` (the reference's equivalent of the
table is the Cassandra PVC, ``helm/values.yaml:25-28``).  Rank 0 builds the :class:`ShardedB200VectorStore` face --
the object the reference's worker calls ``retriever.invoke(q, filter=...)`` on
(``rag_worker/src/worker/services/agent_graph.py:249,268``; one worker process, ``rag-worker-deployment.yaml:8``) --
and hands it to ``--app module:function`` (the worker's own entry point, called as ``function(store)``); ranks
1..G-1 run the store's receive loop until rank 0 closes it.  Without ``--app``, rank 0 answers JSON lines on stdin::

    {"embedding": [...], "k": 5, "filter": {"namespace": "default", "repo": "acme/api"}}
    -> {"hits": [{"id": ..., "score": ..., "metadata": {...}}, ...]}
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys

import numpy as np


def load_sharded_store(path: str, *, embedding=None, group=None, max_batch: int = 1024):
    """Collective: every rank reads its row range of a ``B200VectorStore.save`` file; rank 0 also keeps the ids,
    text and metadata of all rows.  Returns the :class:`ShardedB200VectorStore` (use it on rank 0,
    ``serve_forever()`` elsewhere)."""
    import torch
    import torch.distributed as dist

    from .index import DeviceShard, MetadataDictionary
    from .sharded import shard_range
    from .sharded_store import CudaRankEngine, ShardedB200VectorStore
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    from .vector_store import B200VectorStore
    meta, all_rows, all_codes, ids, texts, metas = B200VectorStore.read_saved(path)
    n = len(ids)
    lo, hi = shard_range(n, world, rank)
    rows16 = np.ascontiguousarray(all_rows[lo:hi])
    codes = np.ascontiguousarray(all_codes[:, lo:hi]) if all_codes.shape[0] else None
    dic = MetadataDictionary(meta["keys"])
    dic.values = [dict(v) for v in meta["values"]]
    # every rank but the last may grow up to its neighbour's first row; the last one up to the 32-bit row limit
    shard = DeviceShard.from_host(rows16, codes, device=torch.cuda.current_device(), row_offset=lo,
                                  capacity=(hi - lo) + (1024 if rank == world - 1 else 0))
    st = ShardedB200VectorStore(CudaRankEngine(shard), dictionary=dic, embedding=embedding,
                                id_fn=lambda r: ids[r], text_fn=lambda r: texts[r], meta_fn=lambda r: metas[r],
                                metric=meta.get("metric", "cosine"), table_name=meta.get("table_name", "embeddings"),
                                scope=meta.get("scope"), multi_valued_keys=meta.get("multi_valued_keys"),
                                max_batch=max_batch, group=group)
    if rank == 0:
        for i, md in enumerate(metas):
            st._post_row(i, md)
    return st


def _answer_lines(store, inp, out) -> None:
    for line in inp:
        line = line.strip()
        if not line:
            continue
        try:
            req = json.loads(line)
            hits = store.similarity_search_with_score_id_by_vector(req["embedding"], k=int(req.get("k", 5)),
                                                                   filter=req.get("filter"))
            out.write(json.dumps({"hits": [{"id": rid, "score": doc.metadata.get("score"), "metadata":
                                            {a: b for a, b in doc.metadata.items() if not a.startswith("_")},
                                            "text": doc.page_content[:1200]} for doc, _, rid in hits]}) + "\n")
        except Exception as e:  # noqa: BLE001 - a bad request must not end the service
            out.write(json.dumps({"error": f"{type(e).__name__}: {e}"}) + "\n")
        out.flush()


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--load", required=True, help="table saved with B200VectorStore.save")
    ap.add_argument("--app", default=None, help="module:function called on rank 0 as function(store)")
    ap.add_argument("--max-batch", type=int, default=1024)
    args = ap.parse_args(argv)
    import torch
    import torch.distributed as dist
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    try:
        store = load_sharded_store(args.load, max_batch=args.max_batch)
        if rank != 0:
            store.serve_forever()
            return 0
        try:
            if args.app:
                mod, fn = args.app.split(":")
                getattr(importlib.import_module(mod), fn)(store)
            else:
                _answer_lines(store, sys.stdin, sys.stdout)
        finally:
            store.close()
        return 0
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())

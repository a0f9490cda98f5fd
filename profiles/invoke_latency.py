"""End-to-end latency of the worker's call -- retriever.invoke(q, filter=...) (agent_graph.py:249) --
through the factory mirror over a 1M x 384 chunk table with the reference's metadata vocabulary:
one start search + one batched adjacent() fan-out per depth."""
import json, sys, time
import numpy as np
sys.path.insert(0, ".")
from githubrepostorag_b200 import synthetic as S
from githubrepostorag_b200.index import MetadataDictionary
from githubrepostorag_b200.retrievers import GraphRetrieverFactory, TableNames
from githubrepostorag_b200.vector_store import B200VectorStore

n, d = 1_000_000, 384
rng = np.random.default_rng(1)
emb16 = S.corpus(1001, n, d)
repo = rng.zipf(1.3, n) % 500
module = (repo * 37 + rng.integers(0, 20, n)) % 6000
fpath = (module * 11 + rng.integers(0, 8, n)) % 60000
keys = ("namespace", "repo", "module", "file_path")
dic = MetadataDictionary(keys)
dic.register_vocab("namespace", ["default"]); dic.register_vocab("repo", [f"repo{i}" for i in range(500)])
dic.register_vocab("module", [f"mod{i}" for i in range(6000)]); dic.register_vocab("file_path", [f"src/f{i}.py" for i in range(60000)])
codes = np.stack([np.zeros(n, np.int32), repo.astype(np.int32), module.astype(np.int32), fpath.astype(np.int32)])

class Emb:
    def embed_query(self, text):
        return (emb16[int(text)].astype(np.float32) + 0.3 * np.random.default_rng(int(text)).standard_normal(d, dtype=np.float32)).tolist()

st = B200VectorStore.from_arrays(emb16, codes, dic, id_fn=S.row_id, embedding=Emb(),
                                 meta_fn=lambda r: {"namespace": "default", "repo": f"repo{repo[r]}", "module": f"mod{module[r]}",
                                                    "file_path": f"src/f{fpath[r]}.py", "scope": "chunk"},
                                 text_fn=lambda r: f"chunk {r}")
f = GraphRetrieverFactory(tables=TableNames("r", "m", "f", "c"), embedding=Emb(), store_loader=lambda t: st)
for name, build in (("for_chunk(k=10,start_k=3,adjacent_k=8,max_depth=2)", f.for_chunk), ("for_repo(k=6,start_k=2)", f.for_repo)):
    r = build()
    lat = []
    for i in range(60):
        t0 = time.perf_counter()
        docs = r.invoke(str(1000 + 37 * i), filter={"namespace": "default"})
        lat.append((time.perf_counter() - t0) * 1e3)
    lat = np.array(lat[10:])
    print(json.dumps({"retriever": name, "docs": len(docs), "p50_ms": round(float(np.median(lat)), 3),
                      "p99_ms": round(float(np.percentile(lat, 99)), 3)}), flush=True)
q = np.asarray(Emb().embed_query("5"), np.float32)[None, :]
lat = []
for i in range(200):
    t0 = time.perf_counter(); st._shard.search_host(q, 10); lat.append((time.perf_counter() - t0) * 1e3)
print(json.dumps({"call": "crb_search_host batch 1 k=10 (1M x 384)", "p50_ms": round(float(np.median(lat[20:])), 4)}))
# C1 (BASELINE.json config 1: 10k x 384, k=5, one query): the reference's own CPU-runnable case, through host buffers
from githubrepostorag_b200.index import DeviceShard
c1 = DeviceShard.from_host(S.corpus(1001, 10_000, d), None, device=0)
lat = []
for i in range(400):
    t0 = time.perf_counter(); c1.search_host(q, 5); lat.append((time.perf_counter() - t0) * 1e6)
print(json.dumps({"call": "C1 10k x 384 k=5 through crb_search_host", "p50_us": round(float(np.median(lat[50:])), 1),
                  "p99_us": round(float(np.percentile(lat[50:], 99)), 1)}))

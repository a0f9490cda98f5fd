"""Where the host-side time of one batch-1 request goes (1.25M x 768 shard = one of 8 shards of the 10M corpus):
device-resident search, crb_search_host, the store face, and the sharded store with a world of one (shared-memory
dispatch + exchange to itself) with its phases timed.  One GPU:  python profiles/e2e_overhead_probe.py"""
import json, os, socket, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
from githubrepostorag_b200.vector_store import B200VectorStore
from githubrepostorag_b200.sharded_store import CudaRankEngine, ShardedB200VectorStore

n, d, k, reps = int(os.environ.get("ROWS", 1_250_000)), 768, 10, 400
with socket.socket() as s:
    s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
torch.cuda.set_device(0)
dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
qh = [np.random.default_rng(i).standard_normal((1, d), dtype=np.float32) for i in range(4)]
qd = [torch.from_numpy(q).cuda() for q in qh]
ws = sh.new_workspace(1, k, 0)
out = (torch.empty((1, k), dtype=torch.float32, device="cuda"), torch.empty((1, k), dtype=torch.int64, device="cuda"),
       torch.empty((1,), dtype=torch.int32, device="cuda"))
res = {"rows": n}
for i in range(20): sh.search_device(qd[i % 4], k, out=out, workspace=ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(reps): sh.search_device(qd[i % 4], k, out=out, workspace=ws)
e1.record(); torch.cuda.synchronize()
res["device_us"] = round(e0.elapsed_time(e1) / reps * 1e3, 1)

def wall(fn):
    for i in range(20): fn(i)
    t0 = time.perf_counter()
    for i in range(reps): fn(i)
    return round((time.perf_counter() - t0) / reps * 1e6, 1)

res["search_host_us"] = wall(lambda i: sh.search_host(qh[i % 4], k))
st1 = B200VectorStore.from_shard(sh, id_fn=lambda r: f"row-{r:09d}")
res["store_search_rows_us"] = wall(lambda i: st1.search_rows(qh[i % 4], k))

class NoClose:
    def __init__(self, s): self._s = s
    def __getattr__(self, a): return getattr(self._s, a)
    def close(self): pass
eng = CudaRankEngine(NoClose(sh), cap_keys=1 << 14)
st = ShardedB200VectorStore(eng, max_batch=64)
res["sharded_world1_search_rows_us"] = wall(lambda i: st.search_rows(qh[i % 4], k))
# phases: time inside engine.search (the ctypes call + numpy allocs) against the whole call
acc = {"engine": 0.0}
orig = eng.search
def timed(*a, **kw):
    t0 = time.perf_counter(); r = orig(*a, **kw); acc["engine"] += time.perf_counter() - t0; return r
eng.search = timed
tot = wall(lambda i: st.search_rows(qh[i % 4], k))
res["sharded_engine_call_us"] = round(acc["engine"] / (reps + 20) * 1e6, 1)
res["sharded_python_outside_engine_us"] = round(tot - acc["engine"] / (reps + 20) * 1e6, 1)
st.close()
print(json.dumps(res))
dist.destroy_process_group()

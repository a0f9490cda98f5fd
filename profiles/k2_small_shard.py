"""Fixed overhead of the tensor-core path on a small shard (1.25M x 768 = one of 8 shards of 10M)."""
import sys, json, time, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, b, k = 1_250_000, 768, 1024, 10
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
qs = [torch.randn((b, d), device="cuda") for _ in range(3)]
ws = sh.new_workspace(b, k, 0)
for i in range(5): sh.search_device(qs[i % 3], k, workspace=ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 50
t0 = time.perf_counter(); e0.record()
for i in range(steps): sh.search_device(qs[i % 3], k, workspace=ws)
e1.record(); torch.cuda.synchronize(); wall = (time.perf_counter() - t0) / steps * 1e3
ms = e0.elapsed_time(e1) / steps
print(json.dumps({"n": n, "b": b, "ms_event": round(ms, 3), "ms_wall": round(wall, 3), "ideal_ms_at_1.2PF": round(2.0*b*n*d/1.2e15*1e3, 3)}))

"""Cost of the device-side fallback: a tensor-core batch in which 0, 1 or many queries cannot be proven."""
import sys, json, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, b, k = 2_000_000, 768, 256, 10
g = torch.Generator(device="cuda").manual_seed(5)
rows = torch.randn((n, d), device="cuda", generator=g).to(torch.float16)
rows[100_000:100_400] = rows[100_000]          # 400 identical rows: a query near them ties at the k-th boundary
sh = DeviceShard.from_device(rows)
ws = sh.new_workspace(b, k, 0)
def run(nbad, algo):
    q = torch.randn((b, d), device="cuda", generator=g)
    q[:nbad] = rows[100_000].float()
    for _ in range(3): sh.search_device(q, k, workspace=ws, algo=algo)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): out = sh.search_device(q, k, workspace=ws, algo=algo)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20, out
for nbad in (0, 1, 4, 32, 256):
    t2, o2 = run(nbad, 2)
    print(json.dumps({"n": n, "batch": b, "unprovable": nbad, "k2_ms": round(t2, 3)}), flush=True)
t1, _ = run(0, 1)
print(json.dumps({"n": n, "batch": b, "k1_only_ms": round(t1, 3)}))

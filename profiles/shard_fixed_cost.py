"""Fixed cost of one batch-1 search on shards of 10M/8 ... 10M rows x 768 (what limits 8-GPU scaling): the streaming
kernel alone against its pro-rata share of the 10M-row launch, k = 10.  One GPU."""
import json, sys, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
d, k = 768, 10
rows_all = torch.randn((10_000_000, d), device="cuda").to(torch.float16)
q = [torch.randn((1, d), device="cuda") for _ in range(4)]
base = None
for n in (10_000_000, 5_000_000, 2_500_000, 1_250_000, 625_000):
    sh = DeviceShard.from_device(rows_all[:n])
    ws = sh.new_workspace(1, k, 0)
    out = (torch.empty((1, k), dtype=torch.float32, device="cuda"), torch.empty((1, k), dtype=torch.int64, device="cuda"),
           torch.empty((1,), dtype=torch.int32, device="cuda"))
    for i in range(20): sh.search_device(q[i % 4], k, out=out, workspace=ws)
    torch.cuda.synchronize()
    reps = 200
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps): sh.search_device(q[i % 4], k, out=out, workspace=ws)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if base is None: base = ms
    print(json.dumps({"rows": n, "ms": round(ms, 4), "pro_rata_ms": round(base * n / 10_000_000, 4),
                      "fixed_us": round((ms - base * n / 10_000_000) * 1e3, 1), "TBps": round(n * d * 2 / ms / 1e9, 2)}), flush=True)
    sh.close()

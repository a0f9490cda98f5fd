"""Host cost of issuing one tensor-core search (ctypes + plan + launches) against its device time."""
import sys, json, time, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
d, k = 768, 10
for n, b in ((37_888, 1024), (1_250_000, 1024), (1_250_000, 64)):
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    sh = DeviceShard.from_device(rows)
    q = torch.randn((b, d), device="cuda")
    ws = sh.new_workspace(b, k, 0)
    for _ in range(5): sh.search_device(q, k, workspace=ws, algo=2)
    torch.cuda.synchronize()
    steps = 100
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(steps): sh.search_device(q, k, workspace=ws, algo=2)
    t_issue = (time.perf_counter() - t0) / steps * 1e3
    e1.record(); torch.cuda.synchronize()
    t_wall = (time.perf_counter() - t0) / steps * 1e3
    print(json.dumps({"n": n, "batch": b, "host_issue_ms": round(t_issue, 4), "device_ms": round(e0.elapsed_time(e1) / steps, 4),
                      "wall_ms": round(t_wall, 4)}), flush=True)
    sh.close(); del rows

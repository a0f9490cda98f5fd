import sys, json, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
d, b, k = 768, 1024, 10
for n in (37_888, 151_552, 303_104, 606_208, 1_250_000, 2_500_000, 5_000_000):
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    sh = DeviceShard.from_device(rows)
    qs = [torch.randn((b, d), device="cuda") for _ in range(3)]
    ws = sh.new_workspace(b, k, 0)
    for i in range(5): sh.search_device(qs[i % 3], k, workspace=ws, algo=2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 30
    e0.record()
    for i in range(steps): sh.search_device(qs[i % 3], k, workspace=ws, algo=2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(json.dumps({"n": n, "tiles": (n + 255) // 256, "ms": round(ms, 3), "us_per_1k_tiles": round(ms * 1e3 / ((n + 255) // 256) * 1000, 1)}), flush=True)
    sh.close(); del rows

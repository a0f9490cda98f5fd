"""Sweep K1 rows-in-flight (CODERAG_B200_K1_ROWS) per dimension: one process per setting."""
import json, os, sys, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
d = int(sys.argv[1]); n = int(sys.argv[2]); k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
qs = [torch.randn((1, d), device="cuda") for _ in range(4)]
ws = sh.new_workspace(1, k, 0)
for i in range(10): sh.search_device(qs[i % 4], k, workspace=ws, algo=1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 100
e0.record()
for i in range(steps): sh.search_device(qs[i % 4], k, workspace=ws, algo=1)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(json.dumps({"d": d, "n": n, "k": k, "R": os.environ.get("CODERAG_B200_K1_ROWS", "default"), "ms": round(ms, 4),
                  "gbs": round(n * d * 2 / ms / 1e6, 1)}), flush=True)

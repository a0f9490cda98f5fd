"""One tensor-core search over a shard of 148 corpus tiles (launch-list target for the fixed cost)."""
import sys, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, b, k = int(sys.argv[1]) if len(sys.argv) > 1 else 37_888, 768, 1024, 10
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
q = torch.randn((b, d), device="cuda")
ws = sh.new_workspace(b, k, 0)
for _ in range(6): sh.search_device(q, k, workspace=ws, algo=2)
torch.cuda.synchronize()
print("ok")

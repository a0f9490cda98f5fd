"""C5 shard (12.5M x 384, B=64): a few searches at the k given on the command line (launch-list target)."""
import sys, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, b, k = 12_500_000, 384, 64, int(sys.argv[1]) if len(sys.argv) > 1 else 100
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
q = torch.randn((b, d), device="cuda")
ws = sh.new_workspace(b, k, 0)
for _ in range(5): sh.search_device(q, k, workspace=ws)
torch.cuda.synchronize()
print("ok")

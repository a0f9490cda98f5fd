"""A few batch-1 streaming searches over 12.5M x 384 (ncu target)."""
import sys, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d = 12_500_000, 384
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
q = torch.randn((1, d), device="cuda")
ws = sh.new_workspace(1, 10, 0)
for _ in range(5): sh.search_device(q, 10, workspace=ws, algo=1)
torch.cuda.synchronize()
print("ok")

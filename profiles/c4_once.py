"""C4 (10M x 1024, 3-key filter at ~1 %, k=20, B=1) a few times, for ncu.  ROWS / REPS / FUSED env knobs."""
import os, sys, numpy as np, torch
sys.path.insert(0, ".")
import bench
from githubrepostorag_b200.index import DeviceShard
n = int(os.environ.get("ROWS", 10_000_000)); d, k = 1024, int(os.environ.get("K", 20))
rows = torch.randn((n, d), device="cuda").to(torch.float16)
codes = bench.device_codes(torch, 0, n, 1004, torch.device("cuda"))
if os.environ.get("WIDE"):
    codes = torch.stack([c.to(torch.int32) for c in codes]).contiguous()
sh = DeviceShard.from_device(rows, codes)
combo = (codes[0].to(torch.int64) * 2000 + codes[1].to(torch.int64)) * 2 + codes[2].to(torch.int64)
counts = torch.bincount(combo, minlength=16000).double()
j = int(torch.argmin((counts / n - 0.01).abs()))
filt = np.array([[j // 4000, (j // 2) % 2000, j % 2]], np.int32)
if os.environ.get("EMPTY"):          # a filter (almost) nothing passes: the scan, the launch and the tail alone
    j = int(torch.argmin(counts + (counts == 0) * 1e18)); filt = np.array([[j // 4000, (j // 2) % 2000, j % 2]], np.int32)
print("selectivity", float(counts[j] / n), "filter", filt.tolist(), "widths", sh.code_widths)
q = torch.randn((1, d), device="cuda")
ws = sh.new_workspace(1, k, 1)
reps = int(os.environ.get("REPS", 6))
for _ in range(reps):
    sh.search_device(q, k, filt=filt, workspace=ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): sh.search_device(q, k, filt=filt, workspace=ws)
e1.record(); torch.cuda.synchronize()
print("ms per search", e0.elapsed_time(e1) / 50)

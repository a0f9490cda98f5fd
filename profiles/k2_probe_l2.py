"""Is K2's corpus streaming limited by DRAM access pattern or by TMA request rate?  Stream a corpus
that fits in L2 (49k x 768 fp16 = 75 MB) vs one that does not, same kernel, batch 128."""
import sys, json, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
for n in (49152, 98304, 1_000_000, 4_000_000):
    d, b, k = 768, 128, 10
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    sh = DeviceShard.from_device(rows)
    q = torch.randn((b, d), device="cuda")
    ws = sh.new_workspace(b, k, 0)
    for _ in range(5): sh.search_device(q, k, algo=2, workspace=ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 40
    e0.record()
    for _ in range(steps): sh.search_device(q, k, algo=2, workspace=ws)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(json.dumps({"n": n, "ms": round(ms, 4), "corpus_gbs": round(n * d * 2 / ms / 1e6, 1)}), flush=True)
    sh.close()

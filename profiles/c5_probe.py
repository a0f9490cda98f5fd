"""BASELINE.json config 5, one GPU's shard: 12.5M x 384 fp16, batch 64, k = 100 and k = 10."""
import sys, json, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, b = 12_500_000, 384, 64
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
q = torch.randn((b, d), device="cuda")
ks = [int(a) for a in sys.argv[1:]] or [10, 100]
for k in ks:
    ws = sh.new_workspace(b, k, 0)
    for algo in ((0,) if len(sys.argv) > 1 else (0, 1)):
        for _ in range(4): sh.search_device(q, k, workspace=ws, algo=algo)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): sh.search_device(q, k, workspace=ws, algo=algo)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(json.dumps({"n": n, "dim": d, "batch": b, "k": k, "algo": algo, "ms": round(ms, 3),
                          "qps": round(b / ms * 1e3), "corpus_TBps": round(n * d * 2 / ms / 1e9, 2)}), flush=True)

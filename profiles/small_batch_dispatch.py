"""Streaming kernel (algo 1) against tensor-core kernel (algo 2) for tiny batches: where should the plan switch?"""
import sys, json, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
for n, d in ((10_000_000, 768), (12_500_000, 384), (1_000_000, 768), (1_000_000, 384), (200_000, 384)):
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    sh = DeviceShard.from_device(rows)
    for b in (1, 2, 3, 4):
        q = torch.randn((b, d), device="cuda")
        ws = sh.new_workspace(b, 10, 0)
        out = {}
        for algo in (1, 2):
            for _ in range(5): sh.search_device(q, 10, workspace=ws, algo=algo)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(30): sh.search_device(q, 10, workspace=ws, algo=algo)
            e1.record(); torch.cuda.synchronize()
            out[f"algo{algo}_ms"] = round(e0.elapsed_time(e1) / 30, 4)
        print(json.dumps({"n": n, "dim": d, "batch": b, **out}), flush=True)
    sh.close(); del rows

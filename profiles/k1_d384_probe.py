"""Streaming kernel on 12.5M x 384, B=1, k=10 under its tuning knobs (in-process env toggles)."""
import os, sys, json, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d = 12_500_000, 384
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
q = torch.randn((1, d), device="cuda")
ws = sh.new_workspace(1, 10, 0)
for knobs in ({}, {"CODERAG_B200_K1_E3": "0"}, {"CODERAG_B200_K1_HALF_PAIRS": "2"}, {"CODERAG_B200_K1_HALF_PAIRS": "4"},
              {"CODERAG_B200_K1_E3": "0", "CODERAG_B200_K1_HALF_PAIRS": "4"}, {"CODERAG_B200_K1_HALF": "0"}, {}):
    for k_, v in knobs.items(): os.environ[k_] = v
    sh.reload_tuning()
    for _ in range(5): sh.search_device(q, 10, workspace=ws, algo=1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30): sh.search_device(q, 10, workspace=ws, algo=1)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 30
    print(json.dumps({"knobs": knobs, "ms": round(ms, 4), "TBps": round(n * d * 2 / ms / 1e9, 2)}), flush=True)
    for k_ in knobs: os.environ.pop(k_, None)

"""Streaming groups of 1 vs 2/4 queries (CODERAG_B200_K1_GROUP), filtered and small-shard cases the tensor-core path does not take."""
import os, sys, json, torch, numpy as np
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
for n, d in ((1_000_000, 384), (100_000, 384), (1_000_000, 768)):
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    codes = torch.from_numpy((np.arange(n) % 10).astype(np.int32)[None, :].copy()).cuda()
    sh = DeviceShard.from_device(rows, codes=codes)
    for b, filt in ((2, True), (4, True), (4, False)):
        if not filt and n >= 131072: continue          # unfiltered big shards go to the tensor-core kernel anyway
        q = torch.randn((b, d), device="cuda")
        f = np.full((b, 1), 3, np.int32) if filt else None
        ws = sh.new_workspace(b, 10, 1)
        out = {}
        for grp in ("1", "2", "4"):
            os.environ["CODERAG_B200_K1_GROUP"] = grp; sh.reload_tuning()
            for _ in range(5): sh.search_device(q, 10, filt=f, workspace=ws, algo=1)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(50): sh.search_device(q, 10, filt=f, workspace=ws, algo=1)
            e1.record(); torch.cuda.synchronize()
            out[f"group{grp}_ms"] = round(e0.elapsed_time(e1) / 50, 4)
        os.environ.pop("CODERAG_B200_K1_GROUP", None); sh.reload_tuning()
        print(json.dumps({"n": n, "dim": d, "batch": b, "filtered_10pct": filt, **out}), flush=True)
    sh.close(); del rows

"""adjacent() fan-out on the device: ONE query vector, 12 filters from a file's handful of rows to the whole
1M x 384 table -- the per-group kernel against the balanced one (in-process env toggle)."""
import os, sys, json, torch, numpy as np
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, k = 1_000_000, 384, 8
rng = np.random.default_rng(3)
rows = torch.randn((n, d), device="cuda").to(torch.float16)
repo = (rng.zipf(1.3, n) % 500).astype(np.int32)
module = ((repo * 37 + rng.integers(0, 20, n)) % 6000).astype(np.int32)
fpath = ((module * 11 + rng.integers(0, 8, n)) % 60000).astype(np.int32)
codes = torch.from_numpy(np.stack([np.zeros(n, np.int32), repo, module, fpath])).cuda()
sh = DeviceShard.from_device(rows, codes=codes)
for name, wants in (("12 edges under namespace=default", [[0, -1, -1, -1], [0, 1, -1, -1], [0, 2, -1, -1], [0, 1, 38, -1], [0, 1, 40, -1], [0, 2, 75, -1],
                                                           [0, 1, 38, 420], [0, 1, 38, 421], [0, 1, 40, 441], [0, 2, 75, 830], [0, 3, -1, -1], [0, 3, 112, -1]]),
                    ("12 edges inside repo=2", [[0, 2, -1, -1], [0, 2, 81, -1], [0, 2, 75, -1], [0, 2, 76, -1], [0, 2, 77, -1], [0, 2, 75, 830],
                                                 [0, 2, 75, 831], [0, 2, 76, 840], [0, 2, 77, 850], [0, 2, 78, -1], [0, 2, 79, -1], [0, 2, 80, -1]])):
    filt = np.asarray(wants, np.int32)
    q = torch.randn((1, d), device="cuda").repeat(len(wants), 1).contiguous()
    ws = sh.new_workspace(len(wants), k, len(wants))
    out = {}
    for bal in ("0", "1"):
        os.environ["CODERAG_B200_K1_BALANCED"] = bal; sh.reload_tuning()
        for _ in range(5): sh.search_device(q, k, filt=filt, workspace=ws)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50): r = sh.search_device(q, k, filt=filt, workspace=ws)
        e1.record(); torch.cuda.synchronize()
        out["balanced" if bal == "1" else "per_group"] = round(e0.elapsed_time(e1) / 50, 4)
    print(json.dumps({"case": name, "n": n, "dim": d, "ms": out}), flush=True)

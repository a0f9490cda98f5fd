"""In-process A/B of tensor-core tuning knobs (the library reads its env knobs on every call), interleaved
and repeated so clock/power drift hits every setting alike.  Usage: k2_ab.py KNOB v1 v2 ... [--n rows] [--b batch] [--k k] [--d dim]"""
import os, sys, json, statistics, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
args = sys.argv[1:]
def opt(name, dflt):
    if name in args:
        i = args.index(name); v = int(args[i + 1]); del args[i:i + 2]; return v
    return dflt
n, b, k, d = opt("--n", 10_000_000), opt("--b", 1024), opt("--k", 10), opt("--d", 768)
reps, steps = opt("--reps", 5), opt("--steps", 6)
knob, values = args[0], args[1:]
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
qs = [torch.randn((b, d), device="cuda") for _ in range(3)]
ws = sh.new_workspace(b, k, 0)
res = {v: [] for v in values}
for rep in range(reps + 1):
    for v in values:
        os.environ[knob] = v
        for i in range(2): sh.search_device(qs[i % 3], k, workspace=ws)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps): sh.search_device(qs[i % 3], k, workspace=ws)
        e1.record(); torch.cuda.synchronize()
        if rep: res[v].append(e0.elapsed_time(e1) / steps)
print(json.dumps({"n": n, "batch": b, "k": k, "dim": d, "knob": knob,
                  "median_ms": {v: round(statistics.median(t), 4) for v, t in res.items()},
                  "min_ms": {v: round(min(t), 4) for v, t in res.items()}}))

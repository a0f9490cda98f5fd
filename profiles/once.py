"""One workload a few times, for ncu captures.  MODE = k1_768 | k1_384 | k2_1024 | balanced | c4 ; REPS (default 4)."""
import os, sys, numpy as np, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
mode, reps = os.environ.get("MODE", "k1_768"), int(os.environ.get("REPS", 4))
if mode in ("k1_768", "k1_384", "k2_1024"):
    n, d = (10_000_000, 768) if mode != "k1_384" else (12_500_000, 384)
    b = 1024 if mode == "k2_1024" else 1
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    sh = DeviceShard.from_device(rows)
    q = torch.randn((b, d), device="cuda")
    ws = sh.new_workspace(b, 10, 0)
    for _ in range(reps): sh.search_device(q, 10, workspace=ws)
elif mode == "balanced":
    n, d, k = 1_000_000, 384, 8
    rng = np.random.default_rng(3)
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    repo = (rng.zipf(1.3, n) % 500).astype(np.int16)
    module = ((repo.astype(np.int64) * 37 + rng.integers(0, 20, n)) % 6000).astype(np.int16)
    cols = [torch.zeros(n, dtype=torch.int8, device="cuda"), torch.from_numpy(repo).cuda(), torch.from_numpy(module).cuda()]
    sh = DeviceShard.from_device(rows, cols)
    wants = [[0, -1, -1], [0, 1, -1], [0, 2, -1], [0, 1, 38], [0, 1, 40], [0, 2, 75], [0, 3, -1], [0, 3, 112], [0, 4, -1], [0, 5, -1], [0, 6, -1], [0, 7, -1]]
    q = torch.randn((1, d), device="cuda").repeat(len(wants), 1).contiguous()
    ws = sh.new_workspace(len(wants), k, len(wants))
    for _ in range(reps): sh.search_device(q, k, filt=np.asarray(wants, np.int32), workspace=ws)
else:
    raise SystemExit("unknown MODE")
torch.cuda.synchronize()
print("done", mode)

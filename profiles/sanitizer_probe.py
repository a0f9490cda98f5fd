"""Small-shape run of every kernel family for compute-sanitizer (one tool per gpurun call)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
rng = np.random.default_rng(0)
n, d = 20_000, 384
rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
codes = np.stack([rng.integers(0, 3, n), rng.integers(0, 40, n)]).astype(np.int32)
sh = DeviceShard.from_host(rows16, codes, device=0)
q = rng.standard_normal((130, d), dtype=np.float32)
filt = np.full((130, 2), -1, np.int32); filt[:60] = [1, -1]; filt[60:64] = [-1, 7]
for algo in (1, 2):
    s, r, v = sh.search_host(q, 10, filt=filt, algo=algo)
s, r, v = sh.search_host(q[:3], 100)
sh.append(rows16[:500], codes[:, :500]); sh.tombstone([1, 2, 3])
s, r, v = sh.search_host(q[:2], 5, metric="ip")
print("ok", int(v.min()))
sh.close()

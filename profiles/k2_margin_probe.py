"""Measured distance between the tensor-core score and the canonical fp32 score against the margin the exactness
proof allows (crb_index_set_k5_stats), per dimension: fp16-exact unit queries (rounding of the accumulation only),
on all-positive-product rows (truncation bias adds up) and Gaussian rows; and Gaussian queries (query rounding
dominates the margin).  One GPU."""
import json, sys, numpy as np, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
from tests.test_gpu_k2 import _unit_queries_exact_in_fp16
for n, d in ((200_000, 384), (200_000, 768), (150_000, 1024), (80_000, 2048), (40_000, 4096)):
    rng = np.random.default_rng(d)
    B, k = 128, 10
    q = _unit_queries_exact_in_fp16(rng, B, d)
    rows = rng.standard_normal((n, d), dtype=np.float32)
    for b in range(B):
        rows[1000 + 40 * b: 1040 + 40 * b] = np.abs(rng.standard_normal((40, d), dtype=np.float32)) * np.sign(q[b]) + 0.3 * rng.standard_normal((40, d), dtype=np.float32)
    sh = DeviceShard.from_host(rows.astype(np.float16), None, device=0)
    stats = torch.zeros((B, 4), dtype=torch.float32, device="cuda")
    out = {"dim": d, "rows": n, "allowance": (d // 16 + 1) * 2.0 ** -22 + (d // 32 + 8) * 2.0 ** -24}
    for name, qq in (("exact_unit_queries", q), ("gaussian_queries", rng.standard_normal((B, d), dtype=np.float32))):
        sh.set_k5_stats(stats)
        sh.search_device(torch.from_numpy(qq).cuda(), k, algo=2)
        torch.cuda.synchronize()
        sh.set_k5_stats(None)
        st = stats.cpu().numpy()
        out[name] = {"max_err": float(st[:, 0].max()), "margin_min": float(st[:, 1].min()), "worst_ratio": float((st[:, 0] / st[:, 1]).max()),
                     "rescored_mean": float(st[:, 2].mean()), "fell_back": int(B - st[:, 3].sum())}
    print(json.dumps(out), flush=True)
    sh.close()

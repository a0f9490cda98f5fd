"""Per-code row ranges on a table whose rows arrive repository by repository (or that was clustered with
B200VectorStore.compact(cluster_by=...)): a repository-scoped search and a traversal step over nested edges, against the
same table in random row order.  1M x 384, k = 10, crb_search_host / crb_search_host_any_of, p50 us."""
import json, sys, time
import numpy as np
sys.path.insert(0, ".")
from githubrepostorag_b200 import synthetic as S
from githubrepostorag_b200.index import DeviceShard

n, d, k = 1_000_000, 384, 10
rng = np.random.default_rng(7)
rows16 = S.corpus(1001, n, d)
repo = rng.zipf(1.3, n) % 500
module = repo * 12 + rng.integers(0, 12, n)
fpath = module * 8 + rng.integers(0, 8, n)
codes = np.stack([np.zeros(n, np.int64), repo, module, fpath]).astype(np.int32)
order = np.lexsort((fpath, module, repo))
q = S.queries(5, 1, d)


def p50(fn, reps=200):
    for _ in range(20):
        fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); t.append((time.perf_counter() - t0) * 1e6)
    return round(float(np.median(t)), 1)


sizes = np.bincount(repo)
picks = {"largest repository": int(np.argmax(sizes)), "a 1 % repository": int(np.argmin(np.abs(sizes - n // 100))),
         "a 0.1 % repository": int(np.argmin(np.abs(sizes - n // 1000)))}
res = {}
for layout, perm in (("random order", None), ("clustered by (repo, module, file_path)", order)):
    r16 = rows16 if perm is None else rows16[perm]
    cds = codes if perm is None else codes[:, perm]
    sh = DeviceShard.from_host(r16, cds, device=0)
    out = {"code_widths": sh.code_widths}
    for name, rp in picks.items():
        f = np.asarray([[0, rp, -1, -1]], np.int32)
        out[f"{name} ({int(sizes[rp])} rows)"] = p50(lambda: sh.search_host(q, k, filt=f))
        i = int(np.flatnonzero(cds[1] == rp)[0])
        alts = np.asarray([[0, rp, -1, -1], [0, -1, cds[2, i], -1], [0, -1, -1, cds[3, i]]], np.int32)
        out[f"{name}: union of repo / module / file edges"] = p50(lambda: sh.search_host_any_of(q[0], alts, k, vectors=True))
    out["unfiltered"] = p50(lambda: sh.search_host(q, k))
    res[layout] = out
    sh.close()
print(json.dumps(res))

"""Where retriever.invoke() spends its time on the real box: the C calls (crb_search_host_vectors: staging, launches,
kernels, result copy) against the Python around them, per traversal step.  Same table as invoke_latency.py."""
import json, sys, time
import numpy as np
sys.argv = [sys.argv[0]]
src = open("profiles/invoke_latency.py").read().split("for name, build in")[0]
exec(src)
sh = st._shard
orig = sh.search_host
log = []
def timed(q, k, **kw):
    t0 = time.perf_counter()
    r = orig(q, k, **kw)
    log.append((time.perf_counter() - t0, np.asarray(q).reshape(-1, d).shape[0], k))
    return r
sh.search_host = timed
for name, build in (("for_chunk", f.for_chunk), ("for_repo", f.for_repo)):
    r = build()
    for i in range(10):
        r.invoke(str(1000 + 37 * i), filter={"namespace": "default"})
    tot, per_step = [], {}
    for i in range(100):
        log.clear()
        t0 = time.perf_counter()
        r.invoke(str(2000 + 37 * i), filter={"namespace": "default"})
        tot.append(time.perf_counter() - t0)
        for j, (t, b, k) in enumerate(log):
            per_step.setdefault(j, []).append((t, b))
    out = {"retriever": name, "invoke_p50_us": round(float(np.median(tot)) * 1e6, 1),
           "c_calls_p50_us": [round(float(np.median([t for t, _ in v])) * 1e6, 1) for _, v in sorted(per_step.items())],
           "batch_p50": [int(np.median([b for _, b in v])) for _, v in sorted(per_step.items())]}
    out["python_us"] = round(out["invoke_p50_us"] - sum(out["c_calls_p50_us"]), 1)
    print(json.dumps(out), flush=True)

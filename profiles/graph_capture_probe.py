"""Can an unfiltered crb_search be captured in a CUDA graph and replayed on new queries?  (No host wait is left on
that path: K1 is one launch; K2 is memset + prep + K2 + K5 + fallback, all stream-ordered.)"""
import sys, json, time, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
n, d, k = 1_000_000, 384, 10
rows = torch.randn((n, d), device="cuda").to(torch.float16)
sh = DeviceShard.from_device(rows)
for b in (1, 8):
    q = torch.randn((b, d), device="cuda")
    ws = sh.new_workspace(b, k, 0)
    scores = torch.empty((b, k), dtype=torch.float32, device="cuda")
    out_rows = torch.empty((b, k), dtype=torch.int64, device="cuda")
    nv = torch.empty((b,), dtype=torch.int32, device="cuda")
    def call():
        return sh.search_device(q, k, workspace=ws, out=(scores, out_rows, nv)) if "out" in sh.search_device.__code__.co_varnames else sh.search_device(q, k, workspace=ws)
    for _ in range(3): ref = call()          # warm-up outside capture: attributes, occupancy caches, tensor maps
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    try:
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                res = call()
        ok = True
    except Exception as e:
        ok = False
        print(json.dumps({"batch": b, "capture": "failed", "error": str(e)[:200]}), flush=True)
        continue
    # replay on NEW query values written into the captured buffer
    q2 = torch.randn((b, d), device="cuda")
    q.copy_(q2)
    g.replay(); torch.cuda.synchronize()
    got = [t.clone() for t in res]
    want = [t.clone() for t in sh.search_device(q2, k)]
    torch.cuda.synchronize()
    same = torch.equal(got[1], want[1]) and torch.equal(got[0].view(torch.int32), want[0].view(torch.int32))
    t0 = time.perf_counter()
    for _ in range(200): g.replay()
    torch.cuda.synchronize(); t_graph = (time.perf_counter() - t0) / 200 * 1e3
    t0 = time.perf_counter()
    for _ in range(200): sh.search_device(q, k, workspace=ws)
    torch.cuda.synchronize(); t_plain = (time.perf_counter() - t0) / 200 * 1e3
    print(json.dumps({"batch": b, "capture": "ok", "replay_matches_fresh_search": bool(same), "ms_graph_replay": round(t_graph, 4), "ms_plain_call": round(t_plain, 4)}), flush=True)

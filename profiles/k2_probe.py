"""K2 probe: time the tensor-core path at several corpus sizes / batches (CUDA events)."""
import sys, json
import torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard

def run(n, d, b, k, steps=8, algo=2):
    rows = torch.randn((n, d), device="cuda", dtype=torch.float32).to(torch.float16)
    sh = DeviceShard.from_device(rows)
    q = [torch.randn((b, d), device="cuda") for _ in range(3)]
    ws = sh.new_workspace(b, k, 0)
    out = None
    for i in range(3):
        out = sh.search_device(q[i % 3], k, algo=algo, workspace=ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        sh.search_device(q[i % 3], k, algo=algo, workspace=ws)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    tf = 2.0 * b * n * d / (ms * 1e-3) / 1e12
    gbs = n * d * 2 / (ms * 1e-3) / 1e9
    print(json.dumps({"n": n, "d": d, "b": b, "k": k, "ms": round(ms, 3), "tflops": round(tf, 1), "corpus_gbs": round(gbs, 1)}), flush=True)
    sh.close(); del rows

if __name__ == "__main__":
    import os
    shapes = [(10_000_000, 768, 1024, 10), (10_000_000, 768, 128, 10), (10_000_000, 768, 8, 10)]
    if os.environ.get("PROBE_B"):
        shapes = [(10_000_000, 768, int(b), 10) for b in os.environ["PROBE_B"].split(",")]
    if os.environ.get("PROBE_FULL"):
        shapes = [(65536, 768, 1024, 10), (1_000_000, 768, 1024, 10), (10_000_000, 768, 1024, 10),
                  (10_000_000, 768, 128, 10), (10_000_000, 768, 8, 10), (10_000_000, 768, 256, 10),
                  (10_000_000, 384, 64, 100), (10_000_000, 1024, 1024, 20)]
    for n, d, b, k in shapes:
        run(n, d, b, k)

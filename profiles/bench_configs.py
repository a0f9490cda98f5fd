"""Measure the other BASELINE.json configs on ONE B200 (CUDA events, queries resident in HBM):
C2 1M x 768 B=1 k=10; C4 10M x 1024, 3-key filter at ~1 % selectivity, k=20, B=1 and B=1024;
C5's per-GPU shard 12.5M x 384, B=64, k=100.  Prints one JSON line per case."""
import json, sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from githubrepostorag_b200 import synthetic as S
from githubrepostorag_b200.index import DeviceShard

HBM = 6548.2


def device_rows(n, d, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    out = torch.empty((n, d), dtype=torch.float16, device="cuda")
    step = 1 << 20
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        out[lo:hi] = torch.randn((hi - lo, d), generator=g, device="cuda").to(torch.float16)
    return out


def timeit(fn, steps, warm=5):
    for i in range(warm): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def case(name, n, d, b, k, steps, codes=None, filt=None, alg_bytes=None, algo=0):
    rows = device_rows(n, d, 7)
    codes_t = None if codes is None else torch.from_numpy(codes).cuda()
    sh = DeviceShard.from_device(rows, codes_t)
    qs = [torch.randn((b, d), device="cuda") for _ in range(4)]
    nf = 0 if filt is None else 1
    ws = sh.new_workspace(b, k, nf)
    f = None if filt is None else np.tile(np.asarray(filt, np.int32), (b, 1))
    ms = timeit(lambda i: sh.search_device(qs[i % 4], k, filt=f, workspace=ws, algo=algo), steps)
    s, r, v = sh.search_device(qs[0], k, filt=f, workspace=ws, algo=algo)
    torch.cuda.synchronize()
    out = {"case": name, "n": n, "d": d, "batch": b, "k": k, "ms_per_batch": round(ms, 4), "qps": round(b / ms * 1e3, 1),
           "n_valid_min": int(v.min().item())}
    if alg_bytes:
        out["algorithmic_gb"] = round(alg_bytes / 1e9, 4)
        out["achieved_gbs"] = round(alg_bytes / ms / 1e6, 1)
        out["frac_hbm_measured"] = round(alg_bytes / ms / 1e6 / HBM, 3)
    else:
        out["tflops"] = round(2.0 * b * n * d / ms / 1e9, 1)
        out["scan_gbs"] = round(n * d * 2 / ms / 1e6, 1)
    print(json.dumps(out), flush=True)
    sh.close(); del rows, sh
    torch.cuda.empty_cache()


if __name__ == "__main__":
    which = sys.argv[1:] or ["C1", "C2", "C4", "C5"]
    if "C1" in which:
        # the reference's own CPU-runnable case: the whole table (7.7 MB) sits in L2, the call is launch latency
        case("C1 10kx384 b1 k5", 10_000, 384, 1, 5, 1000, alg_bytes=10_000 * 384 * 2)
        import time
        rows = device_rows(10_000, 384, 7)
        sh = DeviceShard.from_device(rows)
        qh = np.random.default_rng(0).standard_normal((1, 384), dtype=np.float32)
        for _ in range(50): sh.search_host(qh, 5)
        t0 = time.perf_counter()
        for _ in range(1000): sh.search_host(qh, 5)
        print(json.dumps({"case": "C1 10kx384 b1 k5 through crb_search_host (host buffers in and out)",
                          "ms_per_call": round((time.perf_counter() - t0), 4)}), flush=True)
        sh.close()
    if "C2" in which:
        case("C2 1Mx768 b1 k10", 1_000_000, 768, 1, 10, 400, alg_bytes=1_000_000 * 768 * 2)
    if "C4" in which:
        n, d = 10_000_000, 1024
        codes = S.metadata_codes(S.BASE_SEEDS["C4"], n)
        keys = ("namespace", "repo", "component_kind")
        flt, sel = S.pick_filter(codes, keys, 0.01)
        v = S.vocab()
        mat = np.stack([codes[k] for k in keys]).astype(np.int32)
        want = [v[k].index(flt[k]) for k in keys]
        surv = int(np.all(mat == np.asarray(want)[:, None], axis=0).sum())
        print(json.dumps({"C4_filter": flt, "selectivity": sel, "survivors": surv}), flush=True)
        alg = n * 3 * 4 + surv * d * 2
        case("C4 10Mx1024 filter~1% b1 k20", n, d, 1, 20, 100, codes=mat, filt=want, alg_bytes=alg)
        case("C4 10Mx1024 filter~1% b4 k20", n, d, 4, 20, 100, codes=mat, filt=want, alg_bytes=alg)
        case("C4 10Mx1024 filter~1% b1024 k20 (K2 bitmask full scan)", n, d, 1024, 20, 8, codes=mat, filt=want)
        case("C4 10Mx1024 no filter b1 k20", n, d, 1, 20, 100, alg_bytes=n * d * 2)
    if "C5" in which:
        case("C5 shard 12.5Mx384 b64 k100", 12_500_000, 384, 64, 100, 20)
        case("C5 shard 12.5Mx384 b64 k10", 12_500_000, 384, 64, 10, 20)
        case("C5 shard 12.5Mx384 b1 k100", 12_500_000, 384, 1, 100, 100, alg_bytes=12_500_000 * 384 * 2)

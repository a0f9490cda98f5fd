"""One query under one filter: the fused filter+score kernel against K3 + K1, across selectivities (in-process knob)."""
import json, os, sys, numpy as np, torch
sys.path.insert(0, ".")
from githubrepostorag_b200.index import DeviceShard
for n, d, k in ((10_000_000, 1024, 20), (10_000_000, 768, 10), (1_000_000, 384, 8)):
    rows = torch.randn((n, d), device="cuda").to(torch.float16)
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    cols = [torch.zeros(n, dtype=torch.int8, device="cuda"),
            torch.randint(0, 2, (n,), generator=g, device="cuda").to(torch.int8),
            torch.randint(0, 10, (n,), generator=g, device="cuda").to(torch.int8),
            torch.randint(0, 100, (n,), generator=g, device="cuda").to(torch.int8),
            torch.randint(0, 1000, (n,), generator=g, device="cuda").to(torch.int16)]
    sh = DeviceShard.from_device(rows, cols)
    q = torch.randn((1, d), device="cuda")
    ws = sh.new_workspace(1, k, 1)
    for name, want in (("100%", [0, -1, -1, -1, -1]), ("50%", [-1, 1, -1, -1, -1]), ("10%", [-1, -1, 3, -1, -1]), ("1%", [-1, -1, -1, 7, -1]),
                       ("0.1%", [-1, -1, -1, -1, 5]), ("0.5% 3 keys", [0, 1, -1, 7, -1]),
                       ("0.001%", [-1, -1, -1, 7, 5])):
        filt = np.asarray([want], np.int32)
        out = {}
        for fused in ("1", "0", "auto"):
            if fused == "auto":
                os.environ.pop("CODERAG_B200_K1_FUSED", None)
            else:
                os.environ["CODERAG_B200_K1_FUSED"] = fused
            sh.reload_tuning()
            for _ in range(5): sh.search_device(q, k, filt=filt, workspace=ws)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(40): s_, r_, v_ = sh.search_device(q, k, filt=filt, workspace=ws)
            e1.record(); torch.cuda.synchronize()
            out[{"1": "fused_us", "0": "k3_k1_us", "auto": "default_plan_us"}[fused]] = round(e0.elapsed_time(e1) / 40 * 1e3, 1)
        print(json.dumps({"rows": n, "dim": d, "k": k, "filter": name, **out}), flush=True)
    os.environ.pop("CODERAG_B200_K1_FUSED", None)
    sh.close(); del rows, cols

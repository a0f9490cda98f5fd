#!/usr/bin/env python
"""Retrieval throughput benchmark (BASELINE.json metric: retrieval queries/sec on a 10M x 768
fp16 corpus, k=10, at 1/2/4/8 B200; % of HBM / tensor roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]

A *step* is one search of one batch of B queries over the whole (row-sharded) corpus.  Rank 0
prints ONE JSON line.  `value` is whole-job queries/s with queries resident in HBM, timed with
CUDA events (max over ranks); `e2e` is the same metric through the host-buffer API
(H2D of the queries and D2H of the results inside the timed region); `roofline` is the
dominant kernel against MEASURED_PEAKS.json; `cpu_baseline` is the oracle (numpy restatement of
the reference's CPU path) timed on this box's host cores on a bounded row sample.
`--impl reference` times only that CPU path (rank 0; other ranks exit 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "retrieval_queries_per_sec"
UNIT = "queries/s"
N_ROWS = 10_000_000
DIM = 768
TOPK = 10
SEED = 1003                       # C3 base seed
GEN_BLOCK = 1 << 18
CPU_SAMPLE_ROWS = int(os.environ.get("CODERAG_BENCH_SAMPLE_ROWS", 1_000_000))   # bounded CPU sample (tests shrink it)
N_QUERY_SETS = 4                  # rotate distinct query batches between timed steps


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# ----------------------------------------------------------------------------- synthetic data
def device_corpus(torch, lo: int, hi: int, dim: int, device) -> "torch.Tensor":
    """fp16 rows [lo, hi) generated on the device block by block, seeded per block so every
    rank (and the CPU sample) sees the same row for the same global index."""
    out = torch.empty((hi - lo, dim), dtype=torch.float16, device=device)
    g = torch.Generator(device=device)
    b = lo // GEN_BLOCK
    while b * GEN_BLOCK < hi:
        first = b * GEN_BLOCK
        g.manual_seed(SEED * 1_000_003 + b)
        blk = torch.randn((GEN_BLOCK, dim), generator=g, device=device, dtype=torch.float32).to(torch.float16)
        s, e = max(lo, first) - first, min(hi, first + GEN_BLOCK) - first
        out[first + s - lo:first + e - lo] = blk[s:e]
        b += 1
    return out


def host_queries(batch: int, dim: int, which: int) -> np.ndarray:
    rng = np.random.default_rng([SEED + 1, which])
    return rng.standard_normal((batch, dim), dtype=np.float32)


def host_sample_rows(n: int, dim: int) -> np.ndarray:
    """CPU-only corpus sample for --impl reference (no GPU involved): same shape and
    distribution as the device corpus."""
    out = np.empty((n, dim), dtype=np.float32)
    for lo in range(0, n, GEN_BLOCK):
        hi = min(n, lo + GEN_BLOCK)
        rng = np.random.default_rng([SEED, lo // GEN_BLOCK])
        out[lo:hi] = rng.standard_normal((hi - lo, dim), dtype=np.float32).astype(np.float16)
    return out


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------- CPU baseline
def blas_all_cores():
    """torchrun exports OMP_NUM_THREADS=1; the CPU legs must use every host core they can."""
    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        threadpool_limits(limits=os.cpu_count() or 1)
        return max([i.get("num_threads", 1) for i in threadpool_info()] or [1])
    except Exception:
        return 1


def cpu_baseline(sample32: np.ndarray, n_rows_full: int, dim: int, batch: int, k: int, budget_s: float = 20.0):
    """The oracle's vectorised form (numpy fp32 = the reference CPU path restated) on a row
    sample, scaled linearly in rows to the full corpus."""
    import oracle
    threads = blas_all_cores()
    q = host_queries(max(batch, 1), dim, 0)
    oracle.query_vectorised(q[0], sample32[: 1 << 16], k)     # warm numpy/BLAS
    done, t0 = 0, time.perf_counter()
    while True:
        oracle.query_vectorised(q[done % len(q)], sample32, k)
        done += 1
        el = time.perf_counter() - t0
        if el > budget_s or done >= 64:
            break
    qps_sample = done / el
    scale = len(sample32) / n_rows_full
    one_thread = None
    try:  # BASELINE.md reports the all-core and the single-thread figure
        from threadpoolctl import threadpool_limits
        with threadpool_limits(limits=1):
            t1 = time.perf_counter()
            for i in range(2):
                oracle.query_vectorised(q[i % len(q)], sample32, k)
            one_thread = 2 / (time.perf_counter() - t1) * scale
    except Exception:
        pass
    return {"value": qps_sample * scale, "unit": UNIT, "cores": int(threads), "kind": "port", "value_1_thread": one_thread,
            "sample": f"first {len(sample32):,} of {n_rows_full:,} rows x {dim} (fp32 in RAM), {done} single-query "
                      f"searches in {el:.1f}s = {qps_sample:.3f} q/s on the sample, scaled x{scale:.4g} (linear in rows); "
                      f"host cores {os.cpu_count()}, BLAS threads {threads}"}


def run_reference(args):
    rank, _, world = env_rank()
    if rank != 0:
        return 0
    # bounded sample: keep the whole --steps run within a few minutes
    n_sample = min(CPU_SAMPLE_ROWS if args.steps <= 200 else CPU_SAMPLE_ROWS // 4, N_ROWS)
    sample32 = host_sample_rows(n_sample, DIM)
    import oracle
    threads = blas_all_cores()
    q = host_queries(args.batch, DIM, 0)
    for w in range(args.warmup):
        oracle.query_vectorised(q[w % len(q)], sample32[: 1 << 17], TOPK)
    t0 = time.perf_counter()
    n_q = 0
    for s in range(args.steps):
        # one step = one batch; bounded: at most 2 queries of the batch per step are executed
        for b in range(min(args.batch, 2)):
            oracle.query_vectorised(q[(s + b) % len(q)], sample32, TOPK)
            n_q += 1
    el = time.perf_counter() - t0
    scale = n_sample / N_ROWS
    value = n_q / el * scale
    sample = (f"first {n_sample:,} of {N_ROWS:,} rows x {DIM}, {n_q} single-query searches in {el:.1f}s, "
              f"scaled x{scale:.4g} (linear in rows); host cores {os.cpu_count()}, BLAS threads {threads}")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": el / max(args.steps, 1) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": int(threads), "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)
    return 0


def workload_config(args, world):
    return {"workload": f"{N_ROWS // 1_000_000}Mx{DIM}_fp16_k{TOPK}_b{args.batch}", "n_rows": N_ROWS, "dim": DIM, "k": TOPK,
            "batch": args.batch, "metric": "cosine", "sharding": f"rows/{world}",
            "l2": "inputs larger than L2 (corpus shard >= 1.9 GB vs 126 MB L2); query batches rotate",
            "query_sets": N_QUERY_SETS, "exchange": getattr(args, "exchange_used", "none")}


# ----------------------------------------------------------------------------- our arm
def profile_traffic(kernel: str):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json)."""
    p = ROOT / "profiles" / "traffic.json"
    if p.exists():
        return json.loads(p.read_text()).get(kernel)
    return None


def run_ours(args):
    import torch
    import torch.distributed as dist

    from githubrepostorag_b200 import _native
    from githubrepostorag_b200.index import DeviceShard
    from githubrepostorag_b200.sharded import cuda_searcher, shard_range

    rank, local_rank, world = env_rank()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lo, hi = shard_range(N_ROWS, world, rank)
    n_local = hi - lo
    rows = device_corpus(torch, lo, hi, DIM, dev)
    shard = DeviceShard.from_device(rows, None, row_offset=lo)
    torch.cuda.synchronize()
    searcher, exchange = None, "none"
    if world > 1:
        exchange = args.exchange
        if exchange == "fused":
            try:
                from githubrepostorag_b200.sharded import FusedExchangeSearcher
                searcher = FusedExchangeSearcher(shard, cap_keys=1 << 15)
            except Exception as e:  # noqa: BLE001 - symmetric memory unavailable: NCCL all-gather instead
                if rank == 0:
                    print(f"fused exchange unavailable ({e!r}); using the NCCL all-gather", file=sys.stderr)
                exchange = "nccl"
        if searcher is None:
            searcher = cuda_searcher(shard)
    args.exchange_used = exchange
    k = TOPK
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def measure(B: int, steps: int, warmup: int, sample_clocks: bool):
        q_host = [torch.from_numpy(host_queries(B, DIM, i)).pin_memory() for i in range(N_QUERY_SETS)]
        q_dev = [q.to(dev) for q in q_host]
        out = (torch.empty((B, k), dtype=torch.float32, device=dev), torch.empty((B, k), dtype=torch.int64, device=dev),
               torch.empty((B,), dtype=torch.int32, device=dev))
        ws = shard.new_workspace(B, k, 0)

        def step_device(i):
            q = q_dev[i % N_QUERY_SETS]
            if world > 1:
                return searcher.search(q, k)
            return shard.search_device(q, k, out=out, workspace=ws)

        # ---- device-resident timing (value) ----
        for i in range(warmup):
            step_device(i)
        barrier()
        sampler = ClockSampler(local_rank)
        if rank == 0 and sample_clocks:
            sampler.start()
        launches0 = _native.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for i in range(steps):
            step_device(i)
        ev1.record()
        barrier()
        ms = max_over_ranks(ev0.elapsed_time(ev1))
        launches = _native.launch_count() - launches0
        clocks = sampler.stop() if (rank == 0 and sample_clocks) else None

        # ---- end to end through the host-buffer API (e2e) ----
        res_host = [torch.empty((B, k), dtype=torch.float32).pin_memory(), torch.empty((B, k), dtype=torch.int64).pin_memory(),
                    torch.empty((B,), dtype=torch.int32).pin_memory()]
        res_packed = torch.empty((B * k * 12 + B * 4,), dtype=torch.uint8).pin_memory()
        q_stage = torch.empty((B, DIM), dtype=torch.float32, device=dev)

        def step_e2e(i):
            if world == 1:
                return shard.search_host(q_host[i % N_QUERY_SETS].numpy(), k)
            # serving flow: rank 0 holds the request in pinned memory, everyone searches, rank 0 reads back
            if rank == 0:
                q_stage.copy_(q_host[i % N_QUERY_SETS], non_blocking=True)
            s_, r_, v_, _ = searcher.search(q_stage, k, broadcast=True)
            if rank == 0:
                packed = getattr(searcher, "last_packed", None)
                if packed is not None:                       # scores, rows and n_valid in ONE device-to-host copy
                    res_packed.copy_(packed, non_blocking=True)
                else:
                    res_host[0].copy_(s_, non_blocking=True)
                    res_host[1].copy_(r_, non_blocking=True)
                    res_host[2].copy_(v_, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return res_host

        e2e_steps = max(3, min(steps, 200))
        for i in range(min(warmup, 5)):
            step_e2e(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            step_e2e(i)
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)

        per_step_s = ms / steps * 1e-3
        if B <= 4:
            kern = "k1_stream_kernel"
            bytes_per_launch = n_local * DIM * 2          # algorithmic: fp16 shard read once per batch
            achieved = bytes_per_launch / per_step_s / 1e9
            roof = {"bound": "hbm", "kernel": kern, "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                    "frac": achieved / pk["hbm_gbs"], "traffic": profile_traffic(kern) if world == 1 else None, "peak_source": pk["source"],
                    "algorithmic_bytes_per_launch": bytes_per_launch,
                    "note": "a step is ONE K1 launch per rank (merge fused in its last CTA); duration = CUDA-event step time"}
        else:
            kern = "k2_gemm_topk_pair_kernel" if B > 128 else "k2_gemm_topk_kernel"
            flops = 2.0 * B * n_local * DIM
            achieved = flops / per_step_s / 1e12
            roof = {"bound": "tensor", "kernel": kern, "achieved": achieved, "peak": pk["bf16_tflops_sustained"],
                    "unit": "TFLOP/s", "frac": achieved / pk["bf16_tflops_sustained"],
                    "frac_of_burst_peak": achieved / pk["bf16_tflops"], "traffic": profile_traffic(kern) if world == 1 else None,
                    "peak_source": pk["source"] + " (sustained cuBLAS bf16; burst also given)",
                    "algorithmic_flops_per_launch": flops,
                    "note": "duration = CUDA-event step time (prep + K2 + K5 rescore; K2 is >98 % of it per the ncu launch list)"}
        return {"value": B * steps / (ms * 1e-3), "ms_per_step": ms / steps, "steps": steps, "warmup": warmup,
                "e2e": {"value": B * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": B * DIM * 4,
                        "d2h_bytes_per_step": B * k * 12 + B * 4, "steps": e2e_steps},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof}

    main_m = measure(args.batch, args.steps, args.warmup, True)
    second = None
    if args.batch == 1 and not args.no_secondary:
        second = measure(1024, max(5, min(40, args.steps // 10)), 3, True)

    if rank == 0:
        n_sample = min(CPU_SAMPLE_ROWS, n_local)
        sample32 = rows[:n_sample].to(torch.float32).cpu().numpy()
        base = cpu_baseline(sample32, N_ROWS, DIM, args.batch, k, budget_s=args.cpu_budget)
        line = {"metric": METRIC, "value": main_m["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": main_m["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                "config": workload_config(args, world), "e2e": main_m["e2e"], "gpu_launches": main_m["gpu_launches"],
                "clocks": main_m["clocks"], "roofline": main_m["roofline"], "cpu_baseline": base}
        if second is not None:
            line["batch1024"] = {"workload": f"{N_ROWS // 1_000_000}Mx{DIM}_fp16_k{TOPK}_b1024", "value": second["value"],
                                 "unit": UNIT, "ms_per_step": second["ms_per_step"], "steps": second["steps"],
                                 "e2e": second["e2e"], "gpu_launches": second["gpu_launches"], "clocks": second["clocks"],
                                 "roofline": second["roofline"]}
        emit_line(line)
    barrier()
    shard.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


_JSON_OUT = None


def emit_line(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--exchange", choices=("fused", "nccl"), default="fused",
                    help="N>1: fused NVLink peer-memory exchange (crb_search_exchange) or NCCL all-gather + merge")
    ap.add_argument("--no-secondary", action="store_true", help="skip the batch-1024 (tensor-core) secondary measurement")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    # stdout carries the ONE JSON line and nothing else: libraries that print there (NCCL's version banner does,
    # from C, when NCCL_DEBUG is set) are pointed at stderr for the whole run
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python
"""Retrieval throughput benchmark (BASELINE.json metric: retrieval queries/sec on a 10M x 768 fp16 corpus, k=10, at
1/2/4/8 B200; % of HBM / tensor roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5] [--batch B] [--impl ours|reference]

A *step* is one search of one batch of B queries over the whole (row-sharded) corpus.  Rank 0 prints ONE JSON line.

* `value`      whole-job queries/s with the queries resident in HBM, timed with CUDA events (max over ranks);
* `e2e`        the same metric through the store face (`search_rows` of B200VectorStore / ShardedB200VectorStore):
               host buffers in, host buffers out, H2D and D2H inside the timed region;
* `roofline`   the dominant kernel against MEASURED_PEAKS.json;
* `cpu_baseline`  the oracle (numpy restatement of the reference's CPU path) on this box's host cores, bounded sample;
* `parity_check`  outside the timed region: the shipped path against the oracle's ids on the CPU sample (and, at
               N > 1, fused NVLink exchange == NCCL all-gather variant == planted rows); a mismatch exits non-zero;
* `batch1024`  (default run) the tensor-core leg of the same corpus, median / p99 over >= 20 steps;
* `configs`    (default run) BASELINE.json's other configurations, each with ms and roofline fraction.

`--impl reference` times only the CPU path (rank 0; other ranks exit 0).  Workloads: c2 1M x 768 B=1 k=10; c3 (default)
10M x 768 k=10, B from --batch (1 = the metric's line, 1024 = the tensor-core config); c4 10M x 1024, 3-key filter at
~1 % selectivity, k=20; c5 100M x 384 over 8 GPUs (12.5M rows per GPU on fewer), B=64, k=100.
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
GEN_BLOCK = 1 << 18
CPU_SAMPLE_ROWS = int(os.environ.get("CODERAG_BENCH_SAMPLE_ROWS", 1_000_000))   # bounded CPU sample (tests shrink it)
N_QUERY_SETS = 4                  # rotate distinct query batches between timed steps

WORKLOADS = {
    #        rows        dim   k   batch  seed  filtered
    "c2": dict(n=1_000_000, d=768, k=10, batch=1, seed=1002, filtered=False),
    "c3": dict(n=10_000_000, d=768, k=10, batch=1, seed=1003, filtered=False),
    "c4": dict(n=10_000_000, d=1024, k=20, batch=1, seed=1004, filtered=True),
    "c5": dict(n=100_000_000, d=384, k=100, batch=64, seed=1005, filtered=False),
}


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def workload_spec(name: str, batch, world: int):
    w = dict(WORKLOADS[name])
    if batch:
        w["batch"] = int(batch)
    if name == "c5" and world < 8:
        w["n"] = 12_500_000 * world          # the 100M-row system needs the 8 GPUs; fewer hold 12.5M rows each
    tag = {"c2": "1Mx768", "c3": "10Mx768", "c4": "10Mx1024_filter1pct", "c5": f"{w['n'] // 1_000_000}Mx384"}[name]
    if name == "c5" and w["n"] % 1_000_000:
        tag = f"{w['n'] / 1e6:.1f}Mx384"
    w["name"] = f"{tag}_fp16_k{w['k']}_b{w['batch']}"
    return w


# ----------------------------------------------------------------------------- synthetic data
def sample_rows_host(seed: int, n: int, dim: int) -> np.ndarray:
    """The first `n` rows of the corpus as BOTH arms see them: SURVEY.md §8d's block-seeded numpy generator
    (githubrepostorag_b200/synthetic.py), fp16."""
    from githubrepostorag_b200 import synthetic as S
    return S.corpus(seed, max(n, 1), dim, 0, n)


def device_corpus(torch, lo: int, hi: int, dim: int, seed: int, device, numpy_rows: int):
    """fp16 rows [lo, hi) on the device.  Rows below `numpy_rows` (whole 2**18 blocks) come from the numpy generator
    the CPU arm uses -- identical bits on both arms; the rest are generated on the device block by block, seeded per
    block so every rank sees the same row for the same global index."""
    from githubrepostorag_b200 import synthetic as S
    out = torch.empty((hi - lo, dim), dtype=torch.float16, device=device)
    g = torch.Generator(device=device)
    b = lo // GEN_BLOCK
    while b * GEN_BLOCK < hi:
        first = b * GEN_BLOCK
        s, e = max(lo, first) - first, min(hi, first + GEN_BLOCK) - first
        if first < numpy_rows:
            blk = torch.from_numpy(S.corpus_block(seed, b, dim, GEN_BLOCK)[s:e]).to(device)
        else:
            g.manual_seed(seed * 1_000_003 + b)
            blk = torch.randn((GEN_BLOCK, dim), generator=g, device=device, dtype=torch.float32).to(torch.float16)[s:e]
        out[first + s - lo:first + e - lo] = blk
        b += 1
    return out


def host_queries(seed: int, batch: int, dim: int, which: int) -> np.ndarray:
    rng = np.random.default_rng([seed + 1, which])
    return rng.standard_normal((batch, dim), dtype=np.float32)


def device_codes(torch, lo: int, hi: int, seed: int, device):
    """C4's filterable columns (namespace ~ Zipf over 4, repo ~ Zipf(1.1) over 2000, component_kind 80/20;
    SURVEY.md §8d) for rows [lo, hi), block-seeded like the rows: one column per key, as narrow as the key's
    dictionary allows (int8, int16, int8) -- what a store built from real metadata holds."""
    n = hi - lo
    out = [torch.empty((n,), dtype=dt, device=device) for dt in (torch.int8, torch.int16, torch.int8)]
    p_ns = torch.tensor([1.0 / (i + 1) for i in range(4)], device=device)
    p_repo = 1.0 / torch.arange(1, 2001, device=device, dtype=torch.float32) ** 1.1
    g = torch.Generator(device=device)
    b = lo // GEN_BLOCK
    while b * GEN_BLOCK < hi:
        first = b * GEN_BLOCK
        g.manual_seed((seed + 2) * 1_000_003 + b)
        ns = torch.multinomial(p_ns, GEN_BLOCK, replacement=True, generator=g)
        repo = torch.multinomial(p_repo, GEN_BLOCK, replacement=True, generator=g)
        ck = (torch.rand(GEN_BLOCK, generator=g, device=device) < 0.20).to(torch.int64)
        s, e = max(lo, first) - first, min(hi, first + GEN_BLOCK) - first
        for col, src in zip(out, (ns, repo, ck)):
            col[first + s - lo:first + e - lo] = src[s:e].to(col.dtype)
        b += 1
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


def cpu_baseline(sample32: np.ndarray, w, budget_s: float = 20.0):
    """The oracle's vectorised form (numpy fp32 = the reference CPU path restated) on a row sample, scaled linearly
    in rows to the full corpus."""
    import oracle
    threads = blas_all_cores()
    q = host_queries(w["seed"], max(w["batch"], 1), w["d"], 0)
    oracle.query_vectorised(q[0], sample32[: 1 << 16], w["k"])     # warm numpy/BLAS
    done, t0 = 0, time.perf_counter()
    while True:
        oracle.query_vectorised(q[done % len(q)], sample32, w["k"])
        done += 1
        el = time.perf_counter() - t0
        if el > budget_s or done >= 64:
            break
    qps_sample = done / el
    scale = len(sample32) / w["n"]
    one_thread = None
    try:  # BASELINE.md reports the all-core and the single-thread figure
        from threadpoolctl import threadpool_limits
        with threadpool_limits(limits=1):
            t1 = time.perf_counter()
            for i in range(2):
                oracle.query_vectorised(q[i % len(q)], sample32, w["k"])
            one_thread = 2 / (time.perf_counter() - t1) * scale
    except Exception:
        pass
    return {"value": qps_sample * scale, "unit": UNIT, "cores": int(threads), "kind": "port", "value_1_thread": one_thread,
            "sample": f"first {len(sample32):,} of {w['n']:,} rows x {w['d']} (block-seeded numpy rows, fp32 in RAM), {done} "
                      f"single-query searches in {el:.1f}s = {qps_sample:.3f} q/s on the sample, scaled x{scale:.4g} (linear in "
                      f"rows); host cores {os.cpu_count()}, BLAS threads {threads}"}


def run_reference(args):
    rank, _, world = env_rank()
    if rank != 0:
        return 0
    w = workload_spec(args.workload, args.batch, world)
    # bounded sample: keep the whole --steps run within a few minutes
    n_sample = min(CPU_SAMPLE_ROWS if args.steps <= 200 else CPU_SAMPLE_ROWS // 4, w["n"])
    sample32 = sample_rows_host(w["seed"], n_sample, w["d"]).astype(np.float32)
    import oracle
    threads = blas_all_cores()
    q = host_queries(w["seed"], w["batch"], w["d"], 0)
    for i in range(args.warmup):
        oracle.query_vectorised(q[i % len(q)], sample32[: 1 << 17], w["k"])
    t0 = time.perf_counter()
    n_q = 0
    for s in range(args.steps):
        # one step = one batch; bounded: at most 2 queries of the batch per step are executed
        for b in range(min(w["batch"], 2)):
            oracle.query_vectorised(q[(s + b) % len(q)], sample32, w["k"])
            n_q += 1
    el = time.perf_counter() - t0
    scale = n_sample / w["n"]
    value = n_q / el * scale
    sample = (f"first {n_sample:,} of {w['n']:,} rows x {w['d']} (the block-seeded numpy rows the GPU arm holds as its first "
              f"rows), {n_q} single-query searches in {el:.1f}s, scaled x{scale:.4g} (linear in rows); host cores "
              f"{os.cpu_count()}, BLAS threads {threads}")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": el / max(args.steps, 1) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(w, world, "none"),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": int(threads), "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)
    return 0


def workload_config(w, world, exchange):
    return {"workload": w["name"], "n_rows": w["n"], "dim": w["d"], "k": w["k"],
            "batch": w["batch"], "metric": "cosine", "sharding": f"rows/{world}",
            "l2": "inputs larger than L2 (corpus shard >= 1.5 GB vs 126 MB L2); query batches rotate",
            "query_sets": N_QUERY_SETS, "exchange": exchange}


# ----------------------------------------------------------------------------- our arm
def profile_traffic(kernel: str):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json)."""
    p = ROOT / "profiles" / "traffic.json"
    if p.exists():
        return json.loads(p.read_text()).get(kernel)
    return None


class Bench:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.rank, self.local_rank, self.world = env_rank()
        if self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}: launch with torch.distributed.run")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.pk = peaks()
        self.exchange = "none" if self.world == 1 else args.exchange

    # ---- plumbing ----
    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_ok(self, ok: bool) -> bool:
        t = self.torch.tensor([1.0 if ok else 0.0], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(t.item() > 0.5)

    # ---- corpus of one workload on this rank ----
    def build(self, w, numpy_rows: int):
        from githubrepostorag_b200.index import DeviceShard
        from githubrepostorag_b200.sharded import shard_range
        torch = self.torch
        lo, hi = shard_range(w["n"], self.world, self.rank)
        numpy_rows = -(-numpy_rows // GEN_BLOCK) * GEN_BLOCK if numpy_rows else 0
        rows = device_corpus(torch, lo, hi, w["d"], w["seed"], self.dev, numpy_rows)
        codes = device_codes(torch, lo, hi, w["seed"], self.dev) if w["filtered"] else None
        shard = DeviceShard.from_device(rows, codes, row_offset=lo)
        torch.cuda.synchronize()
        filt, sel = None, None
        if w["filtered"]:
            # the (namespace, repo, component_kind) triple whose joint selectivity is closest to 1 % (whole corpus)
            combo = (codes[0].to(torch.int64) * 2000 + codes[1].to(torch.int64)) * 2 + codes[2].to(torch.int64)
            counts = torch.bincount(combo, minlength=4 * 2000 * 2).to(torch.float64)
            if self.world > 1:
                self.dist.all_reduce(counts)
            j = int(torch.argmin((counts / w["n"] - 0.01).abs()))
            filt = np.array([j // 4000, (j // 2) % 2000, j % 2], dtype=np.int32)
            sel = float(counts[j] / w["n"])
        return {"w": w, "lo": lo, "hi": hi, "rows": rows, "codes": codes, "shard": shard, "filt": filt, "sel": sel}

    def searcher_for(self, shard, kind: str):
        from githubrepostorag_b200.sharded import FusedExchangeSearcher, cuda_searcher
        if kind == "fused":
            return FusedExchangeSearcher(shard, cap_keys=1 << 17, dest_rank=0)
        return cuda_searcher(shard)

    # ---- one measurement: device-resident value, per-step median / p99, clocks, launches ----
    def measure_device(self, c, B: int, steps: int, warmup: int, searcher):
        from githubrepostorag_b200 import _native
        torch, w, shard = self.torch, c["w"], c["shard"]
        k = w["k"]
        q_dev = [torch.from_numpy(host_queries(w["seed"], B, w["d"], i)).to(self.dev) for i in range(N_QUERY_SETS)]
        filt = np.tile(c["filt"], (B, 1)) if c["filt"] is not None else None
        out = (torch.empty((B, k), dtype=torch.float32, device=self.dev), torch.empty((B, k), dtype=torch.int64, device=self.dev),
               torch.empty((B,), dtype=torch.int32, device=self.dev))
        ws = shard.new_workspace(B, k, 1 if filt is not None else 0)

        def step(i):
            q = q_dev[i % N_QUERY_SETS]
            if searcher is not None:
                return searcher.search(q, k, filt=filt) if filt is not None else searcher.search(q, k)
            return shard.search_device(q, k, filt=filt, out=out, workspace=ws)

        for i in range(warmup):
            step(i)
        self.barrier()
        sampler = ClockSampler(self.local_rank)
        if self.rank == 0:
            sampler.start()
        launches0 = _native.launch_count()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        self.barrier()
        evs[0].record()
        for i in range(steps):
            step(i)
            evs[i + 1].record()
        self.barrier()
        ms = self.max_over_ranks(evs[0].elapsed_time(evs[-1]))
        per = np.array([evs[i].elapsed_time(evs[i + 1]) for i in range(steps)])
        launches = _native.launch_count() - launches0
        clocks = sampler.stop() if self.rank == 0 else None
        return {"ms_total": ms, "ms_per_step": ms / steps, "ms_median": self.max_over_ranks(float(np.median(per))),
                "ms_p99": self.max_over_ranks(float(np.percentile(per, 99))), "steps": steps, "warmup": warmup,
                "gpu_launches": int(launches), "clocks": clocks}

    def roofline(self, c, B: int, ms_step: float):
        w, pk = c["w"], self.pk
        n_local = c["hi"] - c["lo"]
        per_step_s = ms_step * 1e-3
        tensor_plan = B >= 2 and n_local >= 131072 and c["filt"] is None
        if c["filt"] is not None:
            surv = c["sel"] * n_local
            widths = sum(c["shard"].code_widths[:3])
            by = n_local * widths + surv * w["d"] * 2
            achieved = by / per_step_s / 1e9
            return {"bound": "hbm", "kernel": "k1_fused_filter_kernel, or k3_filter_kernel + k1_stream_kernel (gather), by estimated selectivity" if B == 1
                    else ("k3_filter_kernel + k1_stream_kernel (gather)" if B <= 2 else "k3_filter + compact + k2 on survivors"),
                    "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": None,
                    "peak_source": pk["source"], "algorithmic_bytes_per_launch": by,
                    "note": f"bytes = rows x {widths} B of code columns (3 keys) + survivors x dim x 2 ({c['sel'] * 100:.3f} % pass); "
                            f"full-scan equivalent {n_local * w['d'] * 2 / per_step_s / 1e9:.0f} GB/s"}
        if not tensor_plan or B < 256:
            kern = "k1_stream_kernel" if not tensor_plan else "k2_gemm_topk_kernel"
            if not tensor_plan and w["d"] <= 512:
                kern = "k1_stream_half_kernel"
            by = n_local * w["d"] * 2          # algorithmic: fp16 shard read once per batch
            achieved = by / per_step_s / 1e9
            return {"bound": "hbm", "kernel": kern, "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                    "frac": achieved / pk["hbm_gbs"], "traffic": profile_traffic(kern) if self.world == 1 and w["n"] == 10_000_000 else None,
                    "peak_source": pk["source"], "algorithmic_bytes_per_launch": by,
                    "note": "duration = CUDA-event step time; B=1: a step is ONE streaming launch per rank (merge fused in its last "
                            "CTA); 2 <= B < 256: prep + tensor-core kernel streaming the shard once + rescore"}
        kern = "k2_gemm_topk_pair_kernel"
        flops = 2.0 * B * n_local * w["d"]
        achieved = flops / per_step_s / 1e12
        return {"bound": "tensor", "kernel": kern, "achieved": achieved, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                "frac": achieved / pk["bf16_tflops"], "frac_of_sustained_peak": achieved / pk["bf16_tflops_sustained"],
                "traffic": profile_traffic(kern) if self.world == 1 and w["n"] == 10_000_000 else None,
                "peak_source": pk["source"] + " (burst cuBLAS bf16: the kernel is timed alone; sustained also given)",
                "algorithmic_flops_per_launch": flops,
                "note": "duration = CUDA-event step time (prep + K2 + K5 rescore + fallback check; K2 is >98 % of it per the ncu launch list)"}

    # ---- end to end through the store face ----
    def measure_e2e(self, c, B: int, steps: int, warmup: int):
        from githubrepostorag_b200.vector_store import B200VectorStore
        w, shard = c["w"], c["shard"]
        k = w["k"]
        q_host = [host_queries(w["seed"], B, w["d"], i) for i in range(N_QUERY_SETS)]
        flt, dic = None, None
        if c["filt"] is not None:
            # the store encodes filter DICTS: give it the vocabulary the synthetic codes stand for
            from githubrepostorag_b200 import synthetic as S
            from githubrepostorag_b200.index import MetadataDictionary
            keys = ("namespace", "repo", "component_kind")
            dic = MetadataDictionary(keys)
            for key in keys:
                dic.register_vocab(key, S.vocab()[key])
            one = {key: S.vocab()[key][int(code)] for key, code in zip(keys, c["filt"])}
            flt = [one] * B
        store = None
        if self.world == 1:
            store = B200VectorStore.from_shard(shard, dic, id_fn=lambda r: f"row-{r:09d}")
        else:
            from githubrepostorag_b200.sharded_store import CudaRankEngine, ShardedB200VectorStore
            store = ShardedB200VectorStore(CudaRankEngine(_NoCloseShard(shard), cap_keys=1 << 17), dictionary=dic,
                                           max_batch=max(B, 64))
        el = 0.0
        if self.rank == 0:
            for i in range(min(warmup, 5)):
                store.search_rows(q_host[i % N_QUERY_SETS], k, flt)
            t0 = time.perf_counter()
            for i in range(steps):
                s_, r_, v_ = store.search_rows(q_host[i % N_QUERY_SETS], k, flt)
            el = time.perf_counter() - t0
            if self.world > 1:
                store.release_peers()
        else:
            store.serve_forever()        # returns when rank 0 releases the peers
        self.barrier()
        el = self.max_over_ranks(el)
        if self.world > 1:
            if self.rank == 0:
                store.close()
            else:
                store.serve_forever()    # ends with rank 0's close()
            self.barrier()
        return {"value": B * steps / el, "unit": UNIT, "h2d_bytes_per_step": B * w["d"] * 4 * self.world,
                "d2h_bytes_per_step": B * k * 12 + B * 4, "steps": steps,
                "through": "B200VectorStore.search_rows" if self.world == 1 else
                           "ShardedB200VectorStore.search_rows (rank 0 owns the face, shared-memory dispatch, NVLink exchange to rank 0)"}

    # ---- parity, outside any timed region ----
    def parity_check(self, c, sample16: np.ndarray):
        """(1) rank 0: the shipped path over a shard of the CPU sample (the corpus's first rows) against the oracle's
        ids; (2) N > 1: one planted query per rank through the fused NVLink exchange and the NCCL variant -- same bits,
        each query's planted row first."""
        import oracle
        from githubrepostorag_b200.index import DeviceShard
        from tests.helpers import assert_topk_parity
        torch, w = self.torch, c["w"]
        k = w["k"]
        n_q, ok, why = 0, True, ""
        try:
            if self.rank == 0 and len(sample16):
                n_s = min(len(sample16), c["hi"] - c["lo"])
                sub = DeviceShard.from_device(c["rows"][:n_s], None if c["codes"] is None else [col[:n_s].contiguous() for col in c["codes"]])
                q = host_queries(w["seed"], 4, w["d"], 99)
                keep = None
                filt = None
                if c["filt"] is not None:
                    hc = [col[:n_s].cpu().numpy() for col in c["codes"]]
                    keep = (hc[0] == c["filt"][0]) & (hc[1] == c["filt"][1]) & (hc[2] == c["filt"][2])
                    filt = np.tile(c["filt"], (1, 1))
                e32 = None
                for plan, qs in (("batch-1", [q[i:i + 1] for i in range(4)]), ("batch-4", [q])):
                    for qb in qs:
                        s_, r_, v_ = sub.search_host(qb, k, filt=None if filt is None else np.tile(c["filt"], (len(qb), 1)))
                        for b in range(len(qb)):
                            s_ref, r_ref = oracle.query_vectorised(qb[b], sample16[:n_s], k, keep=keep)
                            assert int(v_[b]) == len(r_ref), f"{plan}: n_valid {int(v_[b])} vs oracle {len(r_ref)}"
                            if e32 is None:
                                e32 = True
                            assert_topk_parity(r_[b, :len(r_ref)], s_[b, :len(r_ref)], r_ref, s_ref,
                                               lambda row, qv=qb[b]: oracle.score_rows(qv, sample16[row:row + 1].astype(np.float32))[0],
                                               what=f"{w['name']} {plan}")
                            n_q += 1
                sub.close()
            if self.world > 1:
                # one planted query per rank (row lo + 17 of its shard + noise), gathered so every rank searches all of them
                g = torch.Generator(device=self.dev)
                g.manual_seed(4242 + self.rank)
                mine = c["rows"][17].to(torch.float32) + 0.05 * torch.randn(w["d"], generator=g, device=self.dev)
                allq = torch.empty((self.world, w["d"]), dtype=torch.float32, device=self.dev)
                self.dist.all_gather_into_tensor(allq, mine[None, :].contiguous())
                fused = self.searcher_for(c["shard"], "fused")
                fused.exchange.struct.dest_mask = 0           # every rank receives: compare on all of them
                nccl = self.searcher_for(c["shard"], "nccl")
                s1, r1, v1, _ = nccl.search(allq.clone(), k)
                s2, r2, v2, _ = fused.search(allq.clone(), k)
                torch.cuda.synchronize()
                fused.check(v2)
                assert torch.equal(r1, r2) and torch.equal(v1, v2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32)), \
                    "fused NVLink exchange and NCCL all-gather variant disagree"
                from githubrepostorag_b200.sharded import shard_range
                want = [shard_range(w["n"], self.world, r)[0] + 17 for r in range(self.world)]
                assert r2[:, 0].tolist() == want, f"planted rows {want} vs {r2[:, 0].tolist()}"
                n_q += self.world
        except AssertionError as e:
            ok, why = False, str(e)[:300]
        all_ok = self.all_ok(ok)
        return {"ok": all_ok, "queries": n_q, **({"error": why} if why else {}),
                "what": "oracle ids on the CPU sample shard (B=1 and B=4 plans)" +
                        ("; fused exchange == NCCL variant == planted rows on the full sharded corpus" if self.world > 1 else "")}


class _NoCloseShard:
    """The bench reuses one resident shard for several searchers / stores: closing a store must not free it."""

    def __init__(self, shard):
        self._s = shard

    def __getattr__(self, name):
        return getattr(self._s, name)

    def close(self):
        pass


def run_ours(args):
    B_ = Bench(args)
    torch, rank, world = B_.torch, B_.rank, B_.world
    w = workload_spec(args.workload, args.batch, world)
    n_sample = min(CPU_SAMPLE_ROWS, w["n"] // world) if args.workload != "c5" else min(GEN_BLOCK, w["n"] // world)
    c = B_.build(w, n_sample)
    n_local = c["hi"] - c["lo"]
    searcher = B_.searcher_for(c["shard"], B_.exchange) if world > 1 else None

    main_m = B_.measure_device(c, w["batch"], args.steps, args.warmup, searcher)
    e2e = B_.measure_e2e(c, w["batch"], max(3, min(args.steps, 200)), args.warmup)
    roof = B_.roofline(c, w["batch"], main_m["ms_per_step"])

    second = None
    if args.workload == "c3" and w["batch"] == 1 and not args.no_secondary:
        st2 = max(20, min(40, args.steps // 5))
        m2 = B_.measure_device(c, 1024, st2, 3, searcher)
        second = {"workload": w["name"].rsplit("_b", 1)[0] + "_b1024", "value": 1024 * st2 / (m2["ms_total"] * 1e-3), "unit": UNIT,
                  "ms_per_step": m2["ms_per_step"], "ms_median": m2["ms_median"], "ms_p99": m2["ms_p99"], "steps": st2,
                  "e2e": B_.measure_e2e(c, 1024, st2, 3), "gpu_launches": m2["gpu_launches"], "clocks": m2["clocks"],
                  "roofline": B_.roofline(c, 1024, m2["ms_median"])}

    # parity and the CPU baseline need the CPU sample (the corpus's first rows, numpy-generated on both arms)
    sample16 = sample_rows_host(w["seed"], min(n_sample, n_local), w["d"]) if rank == 0 else np.empty((0, w["d"]), np.float16)
    parity = B_.parity_check(c, sample16)

    configs = None
    if args.workload == "c3" and w["batch"] == 1 and not args.no_configs:
        configs = {}
        c["shard"].close()
        del c["rows"], c["shard"], searcher
        c = None
        torch.cuda.empty_cache()
        plan = [("c2", 1, 40), ("c4", 1, 40), ("c4", 1024, 20), ("c5", 64, 20)] if world == 1 else [("c5", 64, 20)]
        built = {}
        for name, batch, st in plan:
            ww = workload_spec(name, batch, world)
            key = f"{name}_b{batch}"
            if name not in built:
                for old in list(built):
                    built.pop(old)["shard"].close()
                torch.cuda.empty_cache()
                built[name] = B_.build(ww, GEN_BLOCK)
            cc = dict(built[name], w=ww)
            srch = B_.searcher_for(cc["shard"], B_.exchange) if world > 1 else None
            m = B_.measure_device(cc, batch, st, 5, srch)
            rf = B_.roofline(cc, batch, m["ms_median"])
            entry = {"workload": ww["name"], "n_rows": ww["n"], "ms_median": m["ms_median"], "ms_p99": m["ms_p99"],
                     "ms_per_step": m["ms_per_step"], "steps": st, "value": batch * st / (m["ms_total"] * 1e-3), "unit": UNIT,
                     "gpu_launches": m["gpu_launches"], "roofline": {kk: rf[kk] for kk in ("bound", "kernel", "achieved", "peak", "unit", "frac", "note")
                                                                      if kk in rf}}
            if cc["sel"] is not None:
                entry["selectivity"] = cc["sel"]
                entry["filter_codes"] = cc["filt"].tolist()
            s16 = sample_rows_host(ww["seed"], GEN_BLOCK, ww["d"]) if rank == 0 else np.empty((0, ww["d"]), np.float16)
            entry["parity_check"] = B_.parity_check(cc, s16)
            configs[key] = entry
        for old in list(built):
            built.pop(old)["shard"].close()

    rc = 0
    if rank == 0:
        base = cpu_baseline(sample16.astype(np.float32), w, budget_s=args.cpu_budget)
        line = {"metric": METRIC, "value": w["batch"] * args.steps / (main_m["ms_total"] * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": main_m["ms_per_step"], "ms_median": main_m["ms_median"],
                "ms_p99": main_m["ms_p99"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16",
                "data": "synthetic", "config": workload_config(w, world, B_.exchange), "e2e": e2e,
                "gpu_launches": main_m["gpu_launches"], "clocks": main_m["clocks"], "roofline": roof, "cpu_baseline": base,
                "parity_check": parity}
        if second is not None:
            line["batch1024"] = second
        if configs is not None:
            line["configs"] = configs
        emit_line(line)
        bad = [parity] + [e["parity_check"] for e in (configs or {}).values()]
        if not all(p["ok"] for p in bad):
            print("PARITY CHECK FAILED: " + json.dumps([p for p in bad if not p["ok"]]), file=sys.stderr)
            rc = 3
    B_.barrier()
    if world > 1:
        B_.dist.destroy_process_group()
    return rc


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
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c3",
                    help="BASELINE.json configuration (default c3 = 10M x 768, k=10; B=1 is the metric's line)")
    ap.add_argument("--batch", type=int, default=0, help="override the workload's batch size (c3: 1024 = the tensor-core config)")
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--exchange", choices=("fused", "nccl"), default="fused",
                    help="N>1: fused NVLink peer-memory exchange (crb_search_exchange) or NCCL all-gather + merge")
    ap.add_argument("--no-secondary", action="store_true", help="skip the batch-1024 (tensor-core) secondary measurement")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE configurations (c2, c4, c5)")
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

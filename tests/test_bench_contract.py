"""CPU check of bench.py's reference arm (the driver runs it at round end): one JSON line with the
contract's keys, produced by the oracle port on a bounded row sample -- no GPU involved."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_reference_arm_prints_one_contract_line():
    env = dict(os.environ, CODERAG_BENCH_SAMPLE_ROWS="20000")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "3"],
                         capture_output=True, text=True, env=env, timeout=300, cwd=str(ROOT))
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "retrieval_queries_per_sec" and d["unit"] == "queries/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["value"] > 0
    assert d["config"]["workload"] == "10Mx768_fp16_k10_b1"
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and "20,000" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", CODERAG_BENCH_SAMPLE_ROWS="20000")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1"],
                         capture_output=True, text=True, env=env, timeout=120, cwd=str(ROOT))
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_reference_arm_runs_every_workload_on_its_own_rows():
    """`--workload c2|c4|c5` select BASELINE.json's other configurations; the reference arm times the oracle on the
    block-seeded numpy rows the GPU arm holds as its first rows, so both arms see identical embeddings."""
    import numpy as np
    sys.path.insert(0, str(ROOT))
    import bench
    from githubrepostorag_b200 import synthetic as S
    names = {"c2": "1Mx768_fp16_k10_b1", "c4": "10Mx1024_filter1pct_fp16_k20_b1", "c5": "100Mx384_fp16_k100_b64"}
    for wl, name in names.items():
        w = bench.workload_spec(wl, 0, 8)
        assert w["name"] == name
        env = dict(os.environ, CODERAG_BENCH_SAMPLE_ROWS="4096")
        out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", wl, "--steps", "1"],
                             capture_output=True, text=True, env=env, timeout=300, cwd=str(ROOT))
        assert out.returncode == 0, out.stderr[-1500:]
        d = json.loads(out.stdout.strip())
        assert d["config"]["workload"] == bench.workload_spec(wl, 0, 1)["name"]        # one process: c5 is one 12.5M-row shard
        assert d["config"]["dim"] == w["d"] and d["config"]["k"] == w["k"]
    assert bench.workload_spec("c5", 0, 2)["name"] == "25Mx384_fp16_k100_b64"        # fewer GPUs hold 12.5M rows each
    assert bench.workload_spec("c3", 1024, 1)["name"] == "10Mx768_fp16_k10_b1024"
    # the sample both arms use is the synthetic module's block-seeded generator, bit for bit
    assert np.array_equal(bench.sample_rows_host(1003, 300, 768), S.corpus(1003, 300, 768))

set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_pytest.log
timeout 200 python profiles/e2e_overhead_probe.py > gpurun_out/r2b_e2e_overhead.json 2> gpurun_out/r2b_e2e_overhead.err; tail -2 gpurun_out/r2b_e2e_overhead.json
ROWS=10000 timeout 100 python profiles/e2e_overhead_probe.py > gpurun_out/r2b_e2e_overhead_10k.json 2>> gpurun_out/r2b_e2e_overhead.err; tail -2 gpurun_out/r2b_e2e_overhead_10k.json
timeout 300 python profiles/invoke_latency.py > gpurun_out/r2b_invoke_latency.jsonl 2> gpurun_out/r2b_invoke_latency.err; cat gpurun_out/r2b_invoke_latency.jsonl
timeout 600 python bench.py > gpurun_out/r2b_bench_n1.json 2> gpurun_out/r2b_bench_n1.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2b_bench_n1.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('parity_check'), d['roofline']['frac'], d.get('clocks'))
PY

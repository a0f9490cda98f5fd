set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
tail -4 gpurun_out/r2e_pytest.log
timeout 600 python bench.py > gpurun_out/r2e_bench_n1.json 2> gpurun_out/r2e_bench_n1.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2e_bench_n1.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('parity_check'), d['roofline']['frac'], d.get('clocks'), d['gpu_launches'])
for k,v in d.get('configs',{}).items(): print(k, v['ms_median'], v['roofline']['frac'], v['parity_check']['ok'], v['gpu_launches'])
PY
tail -3 gpurun_out/r2e_bench_n1.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2

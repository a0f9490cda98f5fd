set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_pytest.log
timeout 300 python profiles/invoke_latency.py > gpurun_out/r2b_invoke_latency.jsonl 2> gpurun_out/r2b_invoke_latency.err; cat gpurun_out/r2b_invoke_latency.jsonl
timeout 250 python profiles/invoke_split.py > gpurun_out/r2b_invoke_split.jsonl 2> gpurun_out/r2b_invoke_split.err; cat gpurun_out/r2b_invoke_split.jsonl
timeout 100 python profiles/invoke_once.py && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2b_invoke_launches.csv python profiles/invoke_once.py > gpurun_out/r2b_invoke_ncu.log 2>&1

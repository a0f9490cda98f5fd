set -x
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
tail -4 gpurun_out/r2d_pytest.log
for g in 0 1 0 1; do
  CODERAG_B200_HOST_GRAPH=$g timeout 120 python profiles/invoke_latency.py 2>/dev/null | tail -3 | sed "s/^/graph=$g /"
done > gpurun_out/r2d_graph_ab.txt
cat gpurun_out/r2d_graph_ab.txt
for g in 0 1; do CODERAG_B200_HOST_GRAPH=$g timeout 100 python profiles/e2e_overhead_probe.py 2>/dev/null | tail -1 | sed "s/^/graph=$g /"; done > gpurun_out/r2d_graph_overhead.txt
cat gpurun_out/r2d_graph_overhead.txt

mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -s -k "traversal_parity or edge_cases" > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log
grep -E "passed|failed|rc=|ties|Error|assert" gpurun_out/r2h_pytest.log | tail -12
for s in c3 c2 c4; do timeout 600 python tools/team_bench.py $s >> gpurun_out/r2h_team_bench.log 2>&1; done
cat gpurun_out/r2h_team_bench.log

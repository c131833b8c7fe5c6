mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -k "full_size_properties_c4 or c5_4k or c1_full" > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
grep -E "passed|failed|rc=|c4 full|C1 1024|Error|assert" gpurun_out/r2g_pytest.log | tail -12
python tools/sweep.py c3 64 stack_shared=0,8,12,16 > gpurun_out/r2g_sweep_c3_stack.log 2>&1
python tools/sweep.py c4 16 stack_shared=0,8,12,16 > gpurun_out/r2g_sweep_c4_stack.log 2>&1
python tools/sweep.py c3 64 node_quorum=4,8,12,16 refill_below=8,16 > gpurun_out/r2g_sweep_c3_quorum.log 2>&1
cut -c1-260 gpurun_out/r2g_sweep_c3_stack.log gpurun_out/r2g_sweep_c4_stack.log gpurun_out/r2g_sweep_c3_quorum.log

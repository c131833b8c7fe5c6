mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s -k "full_size or c5_4k or c1_full or camera or group" > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
grep -E "passed|failed|rc=|c4 full|C1 1024|Error" gpurun_out/r2f_pytest.log | tail -12

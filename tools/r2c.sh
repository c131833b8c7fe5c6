mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "path_kernel or group_api or same_streams" > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -4 gpurun_out/r2c_pytest.log
python tools/overlap_probe.py c3 64 6 > gpurun_out/r2c_overlap.log 2>&1
python tools/overlap_probe.py c3 64 6 extend_blocks=6 nee_blocks=6 >> gpurun_out/r2c_overlap.log 2>&1
python tools/overlap_probe.py c3 64 6 extend_blocks=5 nee_blocks=5 >> gpurun_out/r2c_overlap.log 2>&1
python tools/overlap_probe.py c2 256 6 >> gpurun_out/r2c_overlap.log 2>&1
cat gpurun_out/r2c_overlap.log

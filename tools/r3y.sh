mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "group or two_devices" > gpurun_out/r3y_pytest_2gpu.log 2>&1; tail -2 gpurun_out/r3y_pytest_2gpu.log
bash tools/scale_only.sh 2 r3y

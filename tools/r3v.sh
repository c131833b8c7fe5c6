mkdir -p gpurun_out
rm -f gpurun_out/r3v_*.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tile_lists or multiray or same_streams" > gpurun_out/r3v_pytest.log 2>&1; tail -2 gpurun_out/r3v_pytest.log
export PTB_NOPROF=1
timeout 600 python tools/sweep.py c3 64 tile_nodes=0,1 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3v.log
timeout 600 python tools/sweep.py c4 16 tile_nodes=0,1 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3v.log
cat gpurun_out/r3v.log

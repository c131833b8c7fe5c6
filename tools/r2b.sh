# round-2 step b: k_path correctness + path_threshold sweep
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "path_kernel or group_api or slot_ordered or wavefront_groups or same_streams" > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_pytest.log
python tools/sweep.py c3 64 path_threshold=0,100000,200000,400000,800000,1600000,3200000,6400000 > gpurun_out/r2b_sweep_c3_thr.log 2>&1
python tools/sweep.py c3 64 path_threshold=800000 finish_quorum=2,4,8,16 path_blocks=4,5,6 > gpurun_out/r2b_sweep_c3_fq.log 2>&1
python tools/sweep.py c2 256 path_threshold=0,200000,800000,3200000 > gpurun_out/r2b_sweep_c2_thr.log 2>&1
python tools/sweep.py c4 16 path_threshold=0,200000,800000,3200000 > gpurun_out/r2b_sweep_c4_thr.log 2>&1
cat gpurun_out/r2b_sweep_c3_thr.log | cut -c1-330

mkdir -p gpurun_out
rm -f gpurun_out/r3i_*.log
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "not full_size and not statistical" > gpurun_out/r3i_pytest.log 2>&1; tail -3 gpurun_out/r3i_pytest.log
export PTB_NOPROF=1
for w in "c3 64" "c4 16" "c2 256" "c1 64"; do timeout 600 python tools/sweep.py $w mr_split=1,0 2>&1 | grep -E '^\{' | cut -c1-110 >> gpurun_out/r3i.log; done
cat gpurun_out/r3i.log

mkdir -p gpurun_out
rm -f gpurun_out/r3j_*.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "multiray" > gpurun_out/r3j_pytest.log 2>&1; tail -2 gpurun_out/r3j_pytest.log
timeout 600 python tools/mr_bench.py c3 quick 2>&1 | grep -v '"parity"' >> gpurun_out/r3j_mr.log
timeout 600 python tools/mr_bench.py c4 quick 2>&1 | grep -v '"parity"' >> gpurun_out/r3j_mr.log
cat gpurun_out/r3j_mr.log
export PTB_NOPROF=1
for w in "c3 64" "c4 16"; do timeout 600 python tools/sweep.py $w multiray=2,3 2>&1 | grep -E '^\{' | cut -c1-110 >> gpurun_out/r3j.log; done
cat gpurun_out/r3j.log

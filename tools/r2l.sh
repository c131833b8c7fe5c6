mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2l_noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -3 gpurun_out/r2l_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest_all.log 2>&1; echo "all rc=$?"
tail -3 gpurun_out/r2l_pytest_all.log

mkdir -p gpurun_out
python bench.py > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2e_bench.err
python bench.py --impl reference > gpurun_out/r2e_bench_ref.json 2>> gpurun_out/r2e_bench.err; echo "ref rc=$?"
cut -c1-600 gpurun_out/r2e_bench.json
python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
tail -5 gpurun_out/r2e_pytest.log

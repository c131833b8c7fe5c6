mkdir -p gpurun_out
python bench.py > gpurun_out/r3f_bench.json 2> gpurun_out/r3f_bench.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/r3f_bench_reference_arm.json 2>> gpurun_out/r3f_bench.err; echo "ref rc=$?"
python __graft_entry__.py smoke > gpurun_out/r3f_smoke.log 2>&1; tail -1 gpurun_out/r3f_smoke.log
cut -c1-300 gpurun_out/r3f_bench.json

mkdir -p gpurun_out
rm -f gpurun_out/r2t_mr_bench.log
timeout 900 python tools/mr_bench.py c3 >> gpurun_out/r2t_mr_bench.log 2>&1
timeout 600 python tools/mr_bench.py c4 quick >> gpurun_out/r2t_mr_bench.log 2>&1
timeout 300 python tools/mr_bench.py c2 quick >> gpurun_out/r2t_mr_bench.log 2>&1
cat gpurun_out/r2t_mr_bench.log

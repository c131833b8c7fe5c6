mkdir -p gpurun_out
rm -f gpurun_out/r2u_mr_bench.log
timeout 900 python tools/mr_bench.py c3 >> gpurun_out/r2u_mr_bench.log 2>&1
timeout 600 python tools/mr_bench.py c4 quick >> gpurun_out/r2u_mr_bench.log 2>&1
timeout 300 python tools/mr_bench.py c2 quick >> gpurun_out/r2u_mr_bench.log 2>&1
grep -v '"parity"' gpurun_out/r2u_mr_bench.log | grep -v '"kernel"'
grep -c '"hit_equal": true, "t_bits_equal": true, "uv_bits_equal": true, "prim_diffs": 0, "inst_diffs": 0, "normal_equal": true, "any_equal": true' gpurun_out/r2u_mr_bench.log
grep -c '"parity"' gpurun_out/r2u_mr_bench.log
python tools/mr_ncu_target.py c3 > gpurun_out/r2u_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"trace_wide_kernel|trace_mr_kernel" -f -o gpurun_out/r2u_mr_c3 python tools/mr_ncu_target.py c3 > gpurun_out/r2u_ncu.log 2>&1
cat gpurun_out/r2u_plain.log; tail -3 gpurun_out/r2u_ncu.log

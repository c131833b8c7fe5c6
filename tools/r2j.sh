mkdir -p gpurun_out
python tools/team_ncu_target.py c3 > gpurun_out/r2j_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"trace_wide_kernel|trace_team_kernel" -f -o gpurun_out/r2j_team_c3 python tools/team_ncu_target.py c3 > gpurun_out/r2j_ncu.log 2>&1
tail -3 gpurun_out/r2j_ncu.log

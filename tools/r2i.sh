mkdir -p gpurun_out
rm -f gpurun_out/r2i_team_bench.log
for s in c3 c4; do timeout 600 python tools/team_bench.py $s >> gpurun_out/r2i_team_bench.log 2>&1; done
cat gpurun_out/r2i_team_bench.log

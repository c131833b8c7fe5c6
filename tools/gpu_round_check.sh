# Round check on one B200: GPU tests, smoke, the contract bench line (default workload C3 + sub-records) and the reference arm, the ncu
# launch list of the bench command, and --set full captures of the three wavefront kernels on C2 / C3 / C4.
# usage: bash tools/gpu_round_check.sh <tag>     (outputs gpurun_out/<tag>_*)
T=${1:-r2m}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench.err
python bench.py > gpurun_out/${T}_bench.json 2>> gpurun_out/${T}_bench.err
# launch list of the bench command itself (primary workload only; cold-cache, serialised: compare SHARES)
python bench.py --only --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/${T}_bench_only.json 2>> gpurun_out/${T}_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/${T}_bench_launches.csv python bench.py --only --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/${T}_ncu_bench_launches.log 2>&1
# --set full captures: one wavefront group, so that a captured launch is the launch the bench's roofline pass times (an early, full-size iteration)
for w in "c2 64 15" "c3 64 15" "c4 16 3"; do set -- $w
python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_plain_$1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip $3 --launch-count 3 -f -o gpurun_out/${T}_$1 python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_ncu_$1.log 2>&1
done
tail -3 gpurun_out/${T}_pytest.log; tail -2 gpurun_out/${T}_smoke.log; cut -c1-200 gpurun_out/${T}_bench.json; ls -la gpurun_out

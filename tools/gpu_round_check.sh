# Round-end check on one B200: GPU tests, smoke, the bench lines of every workload, ncu launch lists and --set full captures.
# usage: bash tools/gpu_round_check.sh <tag>     (outputs gpurun_out/<tag>_*)
T=${1:-r1n}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1
python bench.py > gpurun_out/${T}_bench_c2.json 2> gpurun_out/${T}_bench_c2.err
python bench.py --impl reference > gpurun_out/${T}_bench_c2_reference_arm.json 2>> gpurun_out/${T}_bench_c2.err
python bench.py --workload c3 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench_c3.err
python bench.py --workload c4 > gpurun_out/${T}_bench_c4.json 2> gpurun_out/${T}_bench_c4.err
python bench.py --workload c5 --spp 256 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_c5_256spp.json 2> gpurun_out/${T}_bench_c5.err
# launch lists: the default configuration (two wavefront groups); --set full captures: one group, so that a captured launch is the
# launch the bench's roofline pass times (an early, full-size wavefront iteration)
for w in "c2 64 15" "c3 64 15" "c4 16 3"; do set -- $w
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${T}_$1_launches.csv python tools/ncu_target.py $1 $2 > gpurun_out/${T}_ncu_$1_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip $3 --launch-count 3 -f -o gpurun_out/${T}_$1 python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_ncu_$1.log 2>&1
done
tail -3 gpurun_out/${T}_pytest.log; cat gpurun_out/${T}_bench_c*.json | cut -c1-200

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
# with the multiray kernels an iteration is four launches (k_extend_mr, k_extend part 2, k_shade_ordered, k_nee_mr), else three
for w in "c2 64 15 3" "c3 64 16 4" "c4 16 4 4"; do set -- $w
python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_plain_$1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip $3 --launch-count $4 -f -o gpurun_out/${T}_$1 python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_ncu_$1.log 2>&1
# gpurun brings back 64 MiB at most: summarise on the box, keep only the C3 report itself
python tools/ncu_summary.py report gpurun_out/${T}_$1.ncu-rep > gpurun_out/${T}_$1_kernels_ncu.txt 2>&1
python tools/ncu_blocks.py gpurun_out/${T}_$1.ncu-rep k_extend 0.4 > gpurun_out/${T}_$1_extend_sass_blocks.txt 2>&1
[ $1 = c3 ] || rm -f gpurun_out/${T}_$1.ncu-rep
done
python tools/ncu_summary.py launches gpurun_out/${T}_bench_launches.csv > gpurun_out/${T}_bench_launches.txt 2>&1 && rm -f gpurun_out/${T}_bench_launches.csv
tail -3 gpurun_out/${T}_pytest.log; tail -2 gpurun_out/${T}_smoke.log; cut -c1-200 gpurun_out/${T}_bench.json; ls -la gpurun_out

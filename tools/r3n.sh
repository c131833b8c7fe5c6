T=r3n
mkdir -p gpurun_out
python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench.err
python bench.py > gpurun_out/${T}_bench.json 2>> gpurun_out/${T}_bench.err
python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
python bench.py --only --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/${T}_bench_only.json 2>> gpurun_out/${T}_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/${T}_bench_launches.csv python bench.py --only --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/${T}_ncu_bench_launches.log 2>&1
python tools/ncu_summary.py launches gpurun_out/${T}_bench_launches.csv > gpurun_out/${T}_bench_launches.txt 2>&1 && rm -f gpurun_out/${T}_bench_launches.csv
for w in "c3 64 21 5" "c4 16 6 5"; do set -- $w
python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_plain_$1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip $3 --launch-count 4 -f -o gpurun_out/${T}_$1 python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/${T}_ncu_$1.log 2>&1
python tools/ncu_summary.py report gpurun_out/${T}_$1.ncu-rep > gpurun_out/${T}_$1_kernels_ncu.txt 2>&1
python tools/ncu_blocks.py gpurun_out/${T}_$1.ncu-rep '^k_extend$' 0.4 > gpurun_out/${T}_$1_extend_regen_sass_blocks.txt 2>&1
python tools/ncu_blocks.py gpurun_out/${T}_$1.ncu-rep 'k_extend_mr' 0.4 > gpurun_out/${T}_$1_extend_sass_blocks.txt 2>&1
rm -f gpurun_out/${T}_$1.ncu-rep
done
cut -c1-200 gpurun_out/${T}_bench.json

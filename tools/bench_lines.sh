# the contract bench lines of every single-GPU workload + the reference arm: bash tools/bench_lines.sh <tag>
T=${1:-r1o}
mkdir -p gpurun_out
python bench.py > gpurun_out/${T}_bench_c2.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference > gpurun_out/${T}_bench_c2_reference_arm.json 2>> gpurun_out/${T}_bench.err
python bench.py --workload c3 > gpurun_out/${T}_bench_c3.json 2>> gpurun_out/${T}_bench.err
python bench.py --workload c4 > gpurun_out/${T}_bench_c4.json 2>> gpurun_out/${T}_bench.err
cat gpurun_out/${T}_bench_c*.json | cut -c1-160

# bench.py on N GPUs of one box, launched as the driver does (our arm only): bash tools/scale_only.sh <N> <tag>
N=${1:-2}; T=${2:-r3e}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/${T}_bench_${N}gpu.err
cut -c1-250 gpurun_out/${T}_bench_${N}gpu.json

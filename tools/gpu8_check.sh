mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1n_bench_c2_8gpu.json 2> gpurun_out/r1n_bench_c2_8gpu.err
$TR --master-port 29522 bench.py --gpus 8 --workload c5 --spp 2048 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r1n_bench_c5_2048spp_8gpu.json 2> gpurun_out/r1n_bench_c5_8gpu.err
cut -c1-260 gpurun_out/r1n_bench_c2_8gpu.json gpurun_out/r1n_bench_c5_2048spp_8gpu.json; tail -n 3 gpurun_out/r1n_bench_c2_8gpu.err gpurun_out/r1n_bench_c5_8gpu.err

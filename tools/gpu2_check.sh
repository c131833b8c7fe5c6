mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r1m_bench_c2_2gpu.json 2> gpurun_out/r1m_bench_c2_2gpu.err
$TR --master-port 29512 bench.py --gpus 2 --workload c5 --spp 512 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r1m_bench_c5_512spp_2gpu.json 2> gpurun_out/r1m_bench_c5_2gpu.err
$TR --master-port 29513 bench.py --gpus 2 --impl reference --steps 1 --warmup 1 > gpurun_out/r1m_bench_c2_2gpu_reference_arm.json 2>> gpurun_out/r1m_bench_c2_2gpu.err
cut -c1-300 gpurun_out/r1m_bench_c2_2gpu.json gpurun_out/r1m_bench_c5_512spp_2gpu.json gpurun_out/r1m_bench_c2_2gpu_reference_arm.json; tail -3 gpurun_out/r1m_bench_c2_2gpu.err gpurun_out/r1m_bench_c5_2gpu.err

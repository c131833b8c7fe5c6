mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tile_lists or multiray or same_streams" > gpurun_out/r3t_pytest.log 2>&1; tail -2 gpurun_out/r3t_pytest.log
python bench.py --only > gpurun_out/r3t_bench_only.json 2> gpurun_out/r3t.err; python -c "
import json
r=json.loads(open('gpurun_out/r3t_bench_only.json').read().strip().splitlines()[-1]); print(r['value'], r['ms_per_step'], r['roofline']['frac'], r['rays'])"

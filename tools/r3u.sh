mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tile_lists or multiray or same_streams or groups" > gpurun_out/r3u_pytest.log 2>&1; tail -2 gpurun_out/r3u_pytest.log
python bench.py --only --no-cpu-baseline > gpurun_out/r3u_bench_only.json 2> gpurun_out/r3u.err; python -c "
import json
r=json.loads(open('gpurun_out/r3u_bench_only.json').read().strip().splitlines()[-1]); print(r['value'], r['ms_per_step'])"

mkdir -p gpurun_out
rm -f gpurun_out/r3a_*.log
run() { # lib scene spp opts...
  lib=$1; s=$2; spp=$3; shift 3
  PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/quick_bench.py $s $spp "$@" 2>&1 | grep -E '^\{' | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); t = r['timing']; print('$lib', r['scene'][:2], json.dumps(r['opts']), 'ms', round(r['render_ms'],1), 'ext', t['extend'][0], 'shade', t['shade'][0], 'nee', t['nee'][0], 'tail', t['tail'][0], 'rays', r['counters']['extend'], r['counters']['shadow'])
" >> gpurun_out/r3a_ab.log
}
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "multiray or same_streams or groups" > gpurun_out/r3a_pytest.log 2>&1; tail -3 gpurun_out/r3a_pytest.log
for rep in 1 2; do
run base c3 64 mr_split=0
run base c3 64 mr_split=1
run base c4 16 mr_split=0
run base c4 16 mr_split=1
run mb5 c3 64 mr_split=0
run mb6 c3 64 mr_split=0
run mb5 c2 256
run mb6 c2 256
run base c2 256
done
cat gpurun_out/r3a_ab.log

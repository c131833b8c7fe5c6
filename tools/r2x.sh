mkdir -p gpurun_out
rm -f gpurun_out/r2x_*.log
run() { # scene spp opts...
  s=$1; spp=$2; shift 2
  timeout 600 python tools/quick_bench.py $s $spp "$@" 2>&1 | grep -E '^\{' | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); t = r['timing']; print(r['scene'][:2], json.dumps(r['opts']), 'ms', round(r['render_ms'],1), 'ext', t['extend'][0], 'nee', t['nee'][0], 'tail', t['tail'][0])
" >> gpurun_out/r2x_sweep.log
}
run c3 64 multiray=0
run c3 64 multiray=2 mr_min_per_lane=8 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=16 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=32 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=64 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=16 mr_ql=12 mr_qi=12
run c3 64 multiray=2 mr_min_per_lane=16 mr_qn=12 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=16 mr_qn=12 mr_ql=6 mr_qi=6
run c4 16 multiray=0
run c4 16 multiray=2 mr_min_per_lane=8 mr_ql=8 mr_qi=8
run c4 16 multiray=2 mr_min_per_lane=32 mr_ql=8 mr_qi=8
run c4 16 multiray=2 mr_min_per_lane=16 mr_qn=12 mr_ql=8 mr_qi=8
run c1 64 multiray=0
run c1 64 multiray=2 mr_min_per_lane=16 mr_ql=8 mr_qi=8
cat gpurun_out/r2x_sweep.log

mkdir -p gpurun_out
rm -f gpurun_out/r2w_*.log
run() { # scene spp opts...
  s=$1; spp=$2; shift 2
  timeout 600 python tools/quick_bench.py $s $spp "$@" 2>&1 | grep -E '^\{' | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); t = r['timing']; print(r['scene'][:2], json.dumps(r['opts']), 'ms', round(r['render_ms'],1), 'ext', t['extend'][0], 'nee', t['nee'][0], 'tail', t['tail'][0])
" >> gpurun_out/r2w_sweep.log
}
run c3 64 multiray=0
run c3 64 multiray=0
run c3 64 multiray=2
run c3 64 multiray=2 mr_min_per_lane=1
run c3 64 multiray=2 mr_min_per_lane=2
run c3 64 multiray=2 mr_min_per_lane=4
run c3 64 multiray=2 mr_min_per_lane=8
run c3 64 multiray=2 mr_min_per_lane=2 mr_free_quorum=16
run c3 64 multiray=2 mr_min_per_lane=2 mr_free_quorum=32
run c3 64 multiray=2 mr_min_per_lane=2 mr_qn=12 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=2 mr_qn=16 mr_ql=8 mr_qi=8
run c3 64 multiray=2 mr_min_per_lane=2 mr_qn=20 mr_ql=12 mr_qi=12
run c3 64 multiray=2 mr_min_per_lane=2 mr_qn=16 mr_ql=16 mr_qi=16
run c3 64 multiray=2 mr_min_per_lane=2 mr_qn=16 mr_ql=12 mr_qi=6
run c3 64 multiray=2 mr_min_per_lane=2 mr_qn=16 mr_ql=6 mr_qi=12
run c3 64 multiray=2 mr_min_per_lane=2 mr_blocks=7
run c3 64 multiray=2 mr_min_per_lane=2 mr_blocks=6
run c3 64 multiray=2 mr_min_per_lane=2 wavefront_groups=1
run c3 64 multiray=0 wavefront_groups=1
run c2 256 multiray=2 mr_min_per_lane=2 mr_qn=8 mr_ql=4 mr_qi=4
run c2 256 multiray=2 mr_min_per_lane=2 mr_qn=24 mr_ql=16 mr_qi=16
run c4 16 multiray=2 mr_min_per_lane=2
run c4 16 multiray=2 mr_min_per_lane=2 mr_qn=16 mr_ql=8 mr_qi=8
run c4 16 multiray=2 mr_min_per_lane=2 mr_free_quorum=16
cat gpurun_out/r2w_sweep.log

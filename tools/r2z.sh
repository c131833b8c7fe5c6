mkdir -p gpurun_out
rm -f gpurun_out/r2z_*.log
run() { # lib scene spp opts...
  lib=$1; s=$2; spp=$3; shift 3
  PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/quick_bench.py $s $spp "$@" 2>&1 | grep -E '^\{' | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); t = r['timing']; print('$lib', r['scene'][:2], json.dumps(r['opts']), 'ms', round(r['render_ms'],1), 'ext', t['extend'][0], 'shade', t['shade'][0], 'nee', t['nee'][0], 'tail', t['tail'][0])
" >> gpurun_out/r2z_ab.log
}
for rep in 1 2; do
run base c2 256 shade_class=0
run base c2 256
run noinline c2 256
run base c3 64 shade_class=0
run base c3 64
run noinline c3 64
run base c4 16 shade_class=0
run base c4 16
run noinline c4 16
done
cat gpurun_out/r2z_ab.log

mkdir -p gpurun_out
rm -f gpurun_out/r2v_*.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "multiray" > gpurun_out/r2v_pytest.log 2>&1
tail -5 gpurun_out/r2v_pytest.log
for s in c3 c2 c4; do
  spp=64; [ $s = c2 ] && spp=256; [ $s = c4 ] && spp=16
  for mr in 0 2 3; do
    echo "== $s multiray=$mr" >> gpurun_out/r2v_bench.log
    timeout 600 python tools/quick_bench.py $s $spp multiray=$mr 2>&1 | grep -E '^\{' | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(json.dumps(dict(scene=r['scene'], spp=r['spp'], render_ms=round(r['render_ms'],1), mrays_s=round(r['mrays_s'],1), timing=r['timing'], opts=r['opts'])))
" >> gpurun_out/r2v_bench.log
  done
done
cat gpurun_out/r2v_bench.log

#!/bin/bash
# A/B of builds under build/ab/ in one gpurun call: tools/ab.sh "c2 64" "c3 64" ...   (tuning aid)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2 > gpurun_out/ab_pytest.log
: > gpurun_out/ab.log
for rep in 1 2; do
for lib in build/ab/*.so; do
  for w in "$@"; do
    echo "== $lib $w" >> gpurun_out/ab.log
    PTB_LIB=$PWD/$lib python tools/quick_bench.py $w 2>&1 | grep -E "mrays_s|trace" | sed -E 's/.*"render_ms": ([0-9.]+).*"mrays_s": ([0-9.]+).*"timing": (\{[^}]*\}).*/render_ms \1 mrays \2 \3/' >> gpurun_out/ab.log
  done
done
done
cat gpurun_out/ab_pytest.log gpurun_out/ab.log

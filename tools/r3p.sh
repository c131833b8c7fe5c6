mkdir -p gpurun_out
rm -f gpurun_out/r3p_*.log
export PTB_NOPROF=1
for lib in cap7 cap15 cap31; do
  echo "== $lib" >> gpurun_out/r3p.log
  PTB_TILE_STATS=1 PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/ncu_target.py c3 1 2>&1 | grep "tile lists" | head -1 >> gpurun_out/r3p.log
  PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/sweep.py c3 64 tile_lists=1 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3p.log
  PTB_TILE_STATS=1 PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/ncu_target.py c4 1 2>&1 | grep "tile lists" | head -1 >> gpurun_out/r3p.log
  PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/sweep.py c4 16 tile_lists=1 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3p.log
done
cat gpurun_out/r3p.log

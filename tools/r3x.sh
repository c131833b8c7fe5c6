mkdir -p gpurun_out
rm -f gpurun_out/r3x.log
export PTB_NOPROF=1
for lib in n1 n2 n3 n5s32; do
  echo "== $lib" >> gpurun_out/r3x.log
  PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/sweep.py c4 16 tile_nodes=1 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3x.log
  PTB_LIB=$PWD/build/ab/$lib.so timeout 600 python tools/sweep.py c3 64 tile_nodes=1 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3x.log
done
cat gpurun_out/r3x.log

mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for lib in build/ab/*.so; do for thr in -1 8000 16000; do
echo "== $lib thr $thr"; PTB_LIB=$PWD/$lib python tools/sweep.py c3 64 wavefront_groups=1 tail_threshold=$thr 2>&1 | sed -E 's/"trace_closest.*//' | cut -c1-250
done; done

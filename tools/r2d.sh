mkdir -p gpurun_out
python tools/overlap_probe.py c3 64 12 > gpurun_out/r2d_overlap.log 2>&1
python tools/overlap_probe.py c2 256 12 >> gpurun_out/r2d_overlap.log 2>&1
python tools/overlap_probe.py c4 16 8 >> gpurun_out/r2d_overlap.log 2>&1
cat gpurun_out/r2d_overlap.log
python tools/sweep.py c3 64 path_threshold=0 > gpurun_out/r2d_sweep_c3.log 2>&1; cut -c1-300 gpurun_out/r2d_sweep_c3.log
python tools/sweep.py c2 256 path_threshold=0 > gpurun_out/r2d_sweep_c2.log 2>&1; cut -c1-300 gpurun_out/r2d_sweep_c2.log

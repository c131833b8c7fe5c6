mkdir -p gpurun_out
rm -f gpurun_out/r3h_*.log
export PTB_NOPROF=1
timeout 600 python tools/sweep.py c3 64 wavefront_groups=1,2 refill_below=2,4,6,8 2>&1 | grep -E '^\{' | cut -c1-120 >> gpurun_out/r3h.log
timeout 600 python tools/sweep.py c4 16 wavefront_groups=1,2 refill_below=2,4,8 2>&1 | grep -E '^\{' | cut -c1-120 >> gpurun_out/r3h.log
timeout 600 python tools/sweep.py c2 256 wavefront_groups=1,2 refill_below=2,4,8 2>&1 | grep -E '^\{' | cut -c1-120 >> gpurun_out/r3h.log
timeout 600 python tools/sweep.py c2 256 wavefront_groups=2 nee_refill_below=2,4,8,12 2>&1 | grep -E '^\{' | cut -c1-120 >> gpurun_out/r3h.log
cat gpurun_out/r3h.log

mkdir -p gpurun_out
rm -f gpurun_out/r3r_*.log
export PTB_NOPROF=1
timeout 600 python tools/sweep.py c3 64 regen_blocks=5,6,7,8 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3r.log
timeout 600 python tools/sweep.py c4 16 regen_blocks=5,6,7,8 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3r.log
timeout 600 python tools/sweep.py c3 64 refill_below=2,4,8 2>&1 | grep -E '^\{' | cut -c1-100 >> gpurun_out/r3r.log
cat gpurun_out/r3r.log

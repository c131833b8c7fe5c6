mkdir -p gpurun_out
rm -f gpurun_out/r3g_*.log
export PTB_NOPROF=1
for a in "wavefront_groups=1,2,3,4" "slots_per_pixel=8,12,16" "tail_threshold=-1,30000,120000,240000" "refill_below=4,8,12,16" "nee_refill_below=4,8,12,16" "dense_threshold=40,60,80" "mr_free_quorum=16,24,32"; do
  timeout 600 python tools/sweep.py c3 64 $a 2>&1 | grep -E '^\{' | cut -c1-120 >> gpurun_out/r3g_c3.log
done
for a in "wavefront_groups=1,2,3" "slots_per_pixel=8,16" "refill_below=4,8,16" "nee_refill_below=4,8,16"; do
  timeout 600 python tools/sweep.py c4 16 $a 2>&1 | grep -E '^\{' | cut -c1-120 >> gpurun_out/r3g_c4.log
done
cat gpurun_out/r3g_c3.log; cat gpurun_out/r3g_c4.log

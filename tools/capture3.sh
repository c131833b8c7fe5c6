mkdir -p gpurun_out
for w in "c2 64" "c3 64" "c4 16"; do set -- $w
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip 15 --launch-count 3 -f -o gpurun_out/r1m_g1_$1 python tools/ncu_target.py $1 $2 wavefront_groups=1 > gpurun_out/r1m_g1_$1.log 2>&1
done
tail -2 gpurun_out/r1m_g1_c4.log

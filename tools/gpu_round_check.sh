mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1m_pytest.log
python __graft_entry__.py smoke > gpurun_out/r1m_smoke.log 2>&1
python bench.py > gpurun_out/r1m_bench_c2.json 2> gpurun_out/r1m_bench_c2.err
python bench.py --impl reference > gpurun_out/r1m_bench_c2_reference_arm.json 2>> gpurun_out/r1m_bench_c2.err
python bench.py --workload c3 > gpurun_out/r1m_bench_c3.json 2> gpurun_out/r1m_bench_c3.err
python bench.py --workload c4 > gpurun_out/r1m_bench_c4.json 2> gpurun_out/r1m_bench_c4.err
python bench.py --workload c5 --spp 256 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r1m_bench_c5_256spp.json 2> gpurun_out/r1m_bench_c5.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1m_c3_launches.csv python tools/ncu_target.py c3 64 > gpurun_out/r1m_ncu_c3_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip 15 --launch-count 3 -f -o gpurun_out/r1m_c3 python tools/ncu_target.py c3 64 > gpurun_out/r1m_ncu_c3.log 2>&1
tail -3 gpurun_out/r1m_pytest.log; cat gpurun_out/r1m_bench_c*.json | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1m_c2_launches.csv python tools/ncu_target.py c2 64 > gpurun_out/r1m_ncu_c2_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade_ordered|k_nee" --launch-skip 15 --launch-count 3 -f -o gpurun_out/r1m_c2 python tools/ncu_target.py c2 64 > gpurun_out/r1m_ncu_c2.log 2>&1

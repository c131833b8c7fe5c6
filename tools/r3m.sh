mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r3m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3m_pytest.log
tail -3 gpurun_out/r3m_pytest.log

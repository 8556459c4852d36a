mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/r9_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r9_pytest.log | cut -c1-400

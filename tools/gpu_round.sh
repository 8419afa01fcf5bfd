mkdir -p gpurun_out/r01v
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r01v/pytest.log 2>&1; tail -4 gpurun_out/r01v/pytest.log
APOP_HEAVY_W=3000 timeout 1200 python -m pytest tests/test_resample_gpu.py tests/test_sweep_gpu.py tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/r01v/pytest_heavy.log 2>&1; tail -6 gpurun_out/r01v/pytest_heavy.log

mkdir -p gpurun_out/r01p
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r01p/pytest.log 2>&1; tail -30 gpurun_out/r01p/pytest.log

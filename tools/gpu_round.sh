mkdir -p gpurun_out/r01o
timeout 900 python -m pytest tests/test_stat_gpu.py -m gpu -x -q > gpurun_out/r01o/pytest.log 2>&1; tail -30 gpurun_out/r01o/pytest.log

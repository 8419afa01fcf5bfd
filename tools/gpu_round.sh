mkdir -p gpurun_out/r01w
timeout 900 python -m pytest tests/test_sharded_gpu.py tests/test_mvn_gpu.py -m gpu -x -q > gpurun_out/r01w/pytest.log 2>&1; tail -25 gpurun_out/r01w/pytest.log

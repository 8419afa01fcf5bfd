mkdir -p gpurun_out/r01q
timeout 900 python -m pytest tests/test_dists2_gpu.py tests/test_dists_gpu.py -m gpu -x -q > gpurun_out/r01q/pytest.log 2>&1; tail -30 gpurun_out/r01q/pytest.log

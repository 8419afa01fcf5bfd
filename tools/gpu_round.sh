set -x
mkdir -p gpurun_out/r01g
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r01g/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r01g/pytest.log
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/r01g/bench.json 2> gpurun_out/r01g/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01g/bench_ref.json 2> gpurun_out/r01g/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01g/launches_T20.csv python bench.py --T 20 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r01g/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_propagate|k_ancestors|k_fold_agg|k_chain|k_finalize|k_seg_scan|k_expand' --launch-skip 40 --launch-count 14 -o gpurun_out/r01g/full python bench.py --T 12 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r01g/ncu_full.log 2>&1
ls -la gpurun_out/r01g
tail -3 gpurun_out/r01g/pytest.log
cat gpurun_out/r01g/bench.json

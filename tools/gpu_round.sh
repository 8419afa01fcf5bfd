mkdir -p gpurun_out/r01h
timeout 900 python -m pytest tests/test_sweep_gpu.py tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/r01h/pytest.log 2>&1; tail -5 gpurun_out/r01h/pytest.log
for z in 0 4000 8000 11264 16000 24000 32768; do
  APOP_NOISE_PREFETCH_TILES=$z timeout 300 python bench.py --T 200 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01h/bench_z$z.json 2>gpurun_out/r01h/bench_z$z.err
  python - <<PY
import json
d=json.load(open('gpurun_out/r01h/bench_z$z.json'))
print($z, round(d['value']/1e9,2), {k:round(v) for k,v in d['path_roofline']['kernel_avg_us'].items()})
PY
done

mkdir -p gpurun_out/r01m
timeout 600 python -m pytest tests/test_sweep_gpu.py tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/r01m/pytest.log 2>&1; tail -2 gpurun_out/r01m/pytest.log
run() { # name, defs
  APOP_NVCC_DEFS="$2" python -c "from anglican_b200 import build; build.build(force=True)" > /dev/null 2>&1
  timeout 300 python bench.py --T 100 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01m/$1.json 2>gpurun_out/r01m/$1.err
  python - <<PY
import json
d=json.load(open('gpurun_out/r01m/$1.json'))
print('$1', '$2', round(d['value']/1e9,2), round(d['e2e']['value']/1e9,2), {k:round(v) for k,v in d['path_roofline']['kernel_avg_us'].items()})
PY
}
run base ""
run base2 ""

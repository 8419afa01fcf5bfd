mkdir -p gpurun_out/r01s
run() { # name, defs
  APOP_NVCC_DEFS="$2" python -c "from anglican_b200 import build; build.build(force=True)" > /dev/null 2>&1
  timeout 300 python bench.py --T 800 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01s/$1.json 2>gpurun_out/r01s/$1.err
  python - <<PY
import json
d=json.load(open('gpurun_out/r01s/$1.json'))
print('$1', '$2', round(d['value']/1e9,2), round(d['e2e']['value']/1e9,2), d['clocks']['sm_mhz'], d['clocks']['reasons'], {k:round(v) for k,v in d['path_roofline']['kernel_avg_us'].items()})
PY
}
run hw2c3 ""
run hw4c2 "-DAPOP_PROP_HW=4 -DAPOP_PROP_CTAS=2"
run hw2c2 "-DAPOP_PROP_HW=2 -DAPOP_PROP_CTAS=2"
run hw2c3_again ""

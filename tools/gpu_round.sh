N=${NGPU:-8}
mkdir -p gpurun_out/r02b
run() { # name args
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N $2 > gpurun_out/r02b/$1_$N.json 2> gpurun_out/r02b/$1_$N.err
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open('gpurun_out/r02b/$1_$N.json') if l.startswith('{')][-1]
    print('$1', $N, round(d['value']/1e9,3), round(d['e2e']['value']/1e9,3), d['ms_per_step'], d['wall_ms_per_step'], {k:round(v) for k,v in d['path_roofline']['kernel_avg_us'].items()})
except Exception as e:
    print('fail', e); print(open('gpurun_out/r02b/$1_$N.err').read()[-1500:])
PY
}
run f4 "--T 400 --steps 2 --warmup 3 --inflight 4"
run f6 "--T 400 --steps 2 --warmup 3 --inflight 6"

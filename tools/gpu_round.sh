N=${NGPU:-8}
mkdir -p gpurun_out/r01x
run() { # name args
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N $2 > gpurun_out/r01x/$1_$N.json 2> gpurun_out/r01x/$1_$N.err
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open('gpurun_out/r01x/$1_$N.json') if l.startswith('{')][-1]
    print('$1', $N, round(d['value']/1e9,3), round(d['e2e']['value']/1e9,3), d['roofline']['kernel'], round(d['roofline']['frac'],3), {k:round(v) for k,v in d['path_roofline']['kernel_avg_us'].items()})
except Exception as e:
    print('fail', e); print(open('gpurun_out/r01x/$1_$N.err').read()[-1500:])
PY
}
run lgmvn64 "--workload lgmvn64 --T 40 --steps 2 --warmup 2"

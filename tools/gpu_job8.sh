#!/bin/bash
# N-GPU job (gpurun --gpus N): one process per GPU checks, scaling of the headline config,
# several populations in flight, config 4 (IPMCMC over the ranks), config 5
N=${NGPU:-8}
R=${1:-r03s}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
show() { python - <<PY
import json
try:
    d = json.loads(open("$O/$1.json").read().strip().splitlines()[-1])
    if "value" in d:
        one = d.get("one_sweep_in_flight", {}).get("value", 0)
        print("$1", "%.1fG" % (d["value"]/1e9), "1fl %.1fG" % (one/1e9), "e2e", d["e2e"]["value"], (d["e2e"].get("error") or "")[-160:],
              {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d.get("log_z_check"))
    else:
        print("$1", {k: d[k] for k in d if k != "node_log_z"})
except Exception as e:
    print("$1 failed", e); print(open("$O/$1.err").read()[-1200:])
PY
}
timeout 400 python -m pytest tests/test_zz_multiprocess_gpu.py -q -x > $O/pytest_mp.log 2>&1; tail -3 $O/pytest_mp.log
tr 500 29512 bench.py --gpus $N --T 400 --steps 2 --warmup 3 > $O/sv_f2.json 2> $O/sv_f2.err; show sv_f2
tr 300 29514 bench.py --gpus $N --T 200 --steps 2 --warmup 3 --inflight 4 > $O/sv_f4.json 2> $O/sv_f4.err; show sv_f4
tr 300 29515 bench.py --gpus $N --T 200 --steps 2 --warmup 3 --inflight 6 > $O/sv_f6.json 2> $O/sv_f6.err; show sv_f6
tr 600 29516 tools/bench_configs.py c4dist --nodes 32 --T ${C4T:-2000} --steps 2 --warmup 2 > $O/c4dist.json 2> $O/c4dist.err; show c4dist
tr 500 29517 bench.py --gpus $N --workload lgmvn64 --steps 2 --warmup 3 > $O/c5.json 2> $O/c5.err; show c5

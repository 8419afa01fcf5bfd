#!/bin/bash
# N-GPU diagnostic of the e2e leg (which wait times out, host cores), short sweeps
N=${NGPU:-8}
R=${1:-r03u}
O=gpurun_out/$R
mkdir -p $O
nproc > $O/nproc.txt; cat $O/nproc.txt
python anglican_b200/build.py > /dev/null 2>&1
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
show() { python - <<PY
import json, re
try:
    d = json.loads(open("$O/$1.json").read().strip().splitlines()[-1])
    one = d.get("one_sweep_in_flight", {}).get("value", 0)
    print("$1", "%.1fG" % (d["value"]/1e9), "1fl %.1fG" % (one/1e9), "e2e", d["e2e"]["value"], (d["e2e"].get("error") or "")[-200:],
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()})
except Exception as e:
    t = open("$O/$1.err").read()
    print("$1 failed", sorted(set(re.findall(r"line \d+, in run_ours", t))), sorted(set(re.findall(r"spin timeout; wait \d+ on rank \d+", t))))
PY
}
tr 300 29512 bench.py --gpus $N --T 200 --steps 2 --warmup 3 > $O/f2.json 2> $O/f2.err; show f2
grep -o "spin timeout; wait [0-9]* on rank [0-9]*" $O/f2.err | sort | uniq -c
APOP_SPIN_SECONDS=40 tr 400 29513 bench.py --gpus $N --T 200 --steps 2 --warmup 3 > $O/f2_spin40.json 2> $O/f2_spin40.err; show f2_spin40
tr 300 29514 bench.py --gpus $N --T 200 --steps 2 --warmup 3 --inflight 1 > $O/f1.json 2> $O/f1.err; show f1

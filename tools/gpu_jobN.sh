#!/bin/bash
# N-GPU final: the driver's own command line (default bench) + optional extra in-flight run
N=${NGPU:-8}
R=${1:-r03v}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
show() { python - <<PY
import json, re
try:
    d = json.loads(open("$O/$1.json").read().strip().splitlines()[-1])
    one = d.get("one_sweep_in_flight", {}).get("value", 0)
    print("$1", "N=%d" % d["n_gpus"], "%.1fG" % (d["value"]/1e9), "1fl %.1fG" % (one/1e9), "e2e", d["e2e"]["value"], (d["e2e"].get("error") or "")[-200:],
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["log_z_check"], d["clocks"])
except Exception as e:
    t = open("$O/$1.err").read()
    print("$1 failed", e, sorted(set(re.findall(r"line \d+, in run_ours", t))), sorted(set(re.findall(r"spin timeout; wait \d+ on rank \d+", t))), t[-600:])
PY
}
tr 900 29512 bench.py --gpus $N --steps 2 --warmup 3 > $O/default_$N.json 2> $O/default_$N.err; show default_$N
if [ "$EXTRA" = "1" ]; then
  tr 400 29513 bench.py --gpus $N --T 400 --steps 2 --warmup 3 --inflight 3 > $O/f3_$N.json 2> $O/f3_$N.err; show f3_$N
fi

#!/bin/bash
# one-GPU: full GPU suite, default bench (log Z per seed recorded in the line), 3 in flight, config 2
R=${1:-r03h}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; tail -5 $O/pytest.log
show() {
  python - <<PY
import json
try:
    d = json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    one = d.get("one_sweep_in_flight", {}).get("value", 0)
    print("$1", "%.2fG" % (d["value"]/1e9), "1fl %.2fG" % (one/1e9), "e2e %.2fG" % (d["e2e"]["value"]/1e9),
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["clocks"]["sm_mhz"],
          d.get("kalman_check"), d["path_roofline"]["frac"])
except Exception as e:
    print("bench $1 failed", e); print(open("$O/bench_$1.err").read()[-1500:])
PY
}
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; show default
timeout 600 python bench.py --T 400 --steps 3 --warmup 3 --no-microbench --no-cpu-baseline > $O/bench_sv400.json 2> $O/bench_sv400.err; show sv400
timeout 600 python bench.py --T 400 --steps 3 --warmup 3 --inflight 3 --no-microbench --no-cpu-baseline > $O/bench_f3.json 2> $O/bench_f3.err; show f3
timeout 600 python bench.py --workload lgssm1d --steps 3 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err; show c2
timeout 600 python bench.py --workload lgmvn64 --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err; show c5

#!/bin/bash
# one-GPU: full suite, chain trace, configs 2 / 3 after a change of the exact-fold iteration
R=${1:-r03q}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; tail -4 $O/pytest.log
for lg in 27 20; do timeout 200 python tools/dbg_resolve.py $((1 << lg)) > $O/dbg_$lg.txt 2>&1; echo "== 2^$lg"; grep "T=12\|u=0.9\|general fold\|fold cycles" $O/dbg_$lg.txt | cut -c1-260; done
show() {
  python - <<PY
import json
try:
    d = json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    one = d.get("one_sweep_in_flight", {}).get("value", 0)
    print("$1", "%.2fG" % (d["value"]/1e9), "1fl %.2fG" % (one/1e9), "e2e %.2fG" % (d["e2e"]["value"]/1e9),
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["clocks"]["sm_mhz"], d["path_roofline"]["frac"], d["log_z_check"])
except Exception as e:
    print("bench $1 failed", e); print(open("$O/bench_$1.err").read()[-1500:])
PY
}
timeout 600 python bench.py --workload lgssm1d --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err; show c2
timeout 600 python bench.py --T 400 --steps 3 --warmup 3 --no-microbench --no-cpu-baseline > $O/bench_sv400.json 2> $O/bench_sv400.err; show sv400

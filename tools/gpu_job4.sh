#!/bin/bash
# one-GPU job: GPU suite, headline bench (short T) after the Box-Muller revert, launch list at 2^20
R=${1:-r03f}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; tail -8 $O/pytest.log
show() {
  python - <<PY
import json
try:
    d = json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    one = d.get("one_sweep_in_flight", {}).get("value", 0)
    print("$1", "%.2fG" % (d["value"]/1e9), "1fl %.2fG" % (one/1e9), "e2e %.2fG" % (d["e2e"]["value"]/1e9),
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["clocks"]["sm_mhz"],
          d.get("kalman_check"), (d.get("roofline") or {}).get("frac"), d["path_roofline"]["frac"])
    if "resample_reweight" in d: print("   microbench", [(m["particles"], round(m["ms"], 3), round(m["frac"], 3)) for m in d["resample_reweight"]["microbench"]])
except Exception as e:
    print("bench $1 failed", e); print(open("$O/bench_$1.err").read()[-1500:])
PY
}
timeout 600 python bench.py --T 400 --steps 3 --warmup 3 > $O/bench_sv400.json 2> $O/bench_sv400.err; show sv400
timeout 600 python bench.py --workload lgssm1d --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err; show c2
python tools/prof_sweep.py 20 40 > $O/plain20.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 250 -c 110 --csv --log-file $O/launches_2p20.csv python tools/prof_sweep.py 20 40 > $O/ncu20.log 2>&1
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("$O/launches_2p20.csv")) if len(r) > 5 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = r[4].split("(")[0][:40]
    agg.setdefault(name, []).append(float(r[-1]))
for k, v in agg.items():
    print("%-42s n=%3d avg %.1f us" % (k, len(v), sum(v) / len(v) / 1e3 if max(v) > 1000 else sum(v) / len(v)))
PY

#!/bin/bash
# one-GPU: windowed head fold - tests, chain trace for three head lengths, config 2, headline
R=${1:-r03m}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; tail -3 $O/pytest.log
for h in 256 1024 512 128; do
  APOP_NVCC_DEFS="-DAPOP_HEAD=$h" python anglican_b200/build.py --force > /dev/null 2>&1
  for lg in 27 20; do timeout 200 python tools/dbg_resolve.py $((1 << lg)) > $O/dbg_${h}_$lg.txt 2>&1; echo "== HEAD $h 2^$lg"; grep "T=12\|general fold\|fold cycles" $O/dbg_${h}_$lg.txt | cut -c1-260; done
done
python anglican_b200/build.py --force > /dev/null 2>&1
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
timeout 300 python tools/prof_columns.py 24 > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv -k regex:k_observe --log-file $O/observe.csv python tools/prof_columns.py 24 > $O/ncu.log 2>&1
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("$O/observe.csv")) if len(r) > 8 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    agg.setdefault(r[4].split("(")[0].replace("void ", "")[:40], {}).setdefault(r[-3], []).append((r[-2], float(r[-1].replace(",", ""))))
sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3}
for k, d in agg.items():
    g = lambda key: (d[key][-1][1] * sc.get(d[key][-1][0], 1)) if key in d else 0
    t = g("gpu__time_duration.sum"); b = g("dram__bytes_read.sum") + g("dram__bytes_write.sum")
    print("%-42s %8.1f us %8.1f MB %6.0f GB/s" % (k, t * 1e6, b / 1e6, b / t / 1e9 if t else 0))
PY

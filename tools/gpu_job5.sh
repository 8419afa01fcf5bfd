#!/bin/bash
# one-GPU final job: adaptive tests, full-length default bench (records log Z per seed),
# reference arm, ncu launch list + --set full of the shipped streaming kernels at 2^27
R=${1:-r03g}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
timeout 600 python -m pytest tests/test_zz_adaptive_gpu.py tests/test_zz_graph_gpu.py -q > $O/pytest_adaptive.log 2>&1; tail -4 $O/pytest_adaptive.log
timeout 900 python bench.py --record-logz > $O/bench_default.json 2> $O/bench_default.err
python - <<PY
import json
try:
    d = json.loads(open("$O/bench_default.json").read().strip().splitlines()[-1])
    print("default", "%.2fG" % (d["value"]/1e9), "1fl %.2fG" % (d["one_sweep_in_flight"]["value"]/1e9), "e2e %.2fG" % (d["e2e"]["value"]/1e9),
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["clocks"], d["path_roofline"]["frac"], d["cpu_baseline"]["value"], d["log_z_check"])
except Exception as e:
    print("bench failed", e); print(open("$O/bench_default.err").read()[-1500:])
PY
timeout 600 python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; tail -c 400 $O/bench_reference.json; echo
python tools/prof_sweep.py 27 6 > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 22 -c 44 --csv --log-file $O/launches_2p27.csv python tools/prof_sweep.py 27 6 > $O/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source off -k 'regex:k_propagate|k_fold_agg|k_ancestors$' -s 6 -c 3 \
    -o $O/prof python tools/prof_sweep.py 27 6 > $O/ncu.log 2>&1
ncu -i $O/prof.ncu-rep --page raw --csv > $O/raw.csv 2> /dev/null
python tools/ncu_summary.py $O/raw.csv > $O/summary.csv; cut -c1-300 $O/summary.csv
rm -f $O/prof.ncu-rep

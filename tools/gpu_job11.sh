#!/bin/bash
# one-GPU, what the driver runs at round end: GPU suite (-x), smoke(), default bench, reference arm
R=${1:-r03x}
O=gpurun_out/$R
mkdir -p $O
python -c "import __graft_entry__ as g; g.build()" > $O/build.log 2>&1; tail -1 $O/build.log
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest.log 2>&1; tail -3 $O/pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 1200 python bench.py > $O/bench_default.json 2> $O/bench_default.err
python - <<PY
import json
try:
    d = json.loads(open("$O/bench_default.json").read().strip().splitlines()[-1])
    print("default", "%.2fG" % (d["value"]/1e9), "1fl %.2fG" % (d["one_sweep_in_flight"]["value"]/1e9), "e2e %.2fG" % (d["e2e"]["value"]/1e9),
          {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["clocks"], d["path_roofline"]["frac"], d["log_z_check"], d["roofline"]["frac"], d["roofline"]["traffic"])
    print([(m["particles"], round(m["ms"], 3), round(m["frac"], 3)) for m in d["resample_reweight"]["microbench"]])
except Exception as e:
    print("bench failed", e); print(open("$O/bench_default.err").read()[-1500:])
PY
timeout 600 python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; tail -c 300 $O/bench_reference.json; echo

#!/bin/bash
# one-GPU check of the current tree: GPU suite, short bench, chain phase timings
R=${1:-r03x}
O=gpurun_out/$R
mkdir -p $O
timeout 700 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; tail -5 $O/pytest.log
timeout 600 python bench.py --T 400 --steps 3 --warmup 3 > $O/bench_T400.json 2> $O/bench_T400.err
python - <<PY
import json
try:
    d = json.loads(open("$O/bench_T400.json").read().strip().splitlines()[-1])
    print(d["value"], d["e2e"]["value"], d["path_roofline"]["kernel_avg_us"], d["clocks"])
except Exception as e:
    print("bench failed", e); print(open("$O/bench_T400.err").read()[-2000:])
PY
timeout 200 python tools/dbg_resolve.py > $O/dbg_resolve.txt 2>&1; cat $O/dbg_resolve.txt

#!/bin/bash
# one-GPU check of the current tree + k_propagate table variants (rebuilt on the box)
R=${1:-r03x}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; tail -5 $O/pytest.log
bench() {
  timeout 600 python bench.py --T 200 --steps 2 --warmup 3 --no-microbench --no-cpu-baseline > $O/bench_$1.json 2> $O/bench_$1.err
  python - <<PY
import json
try:
    d = json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    print("$1", "%.2fG" % (d["value"]/1e9), "1fl %.2fG" % (d["one_sweep_in_flight"]["value"]/1e9), {k: round(v) for k, v in d["path_roofline"]["kernel_avg_us"].items()}, d["clocks"]["sm_mhz"])
except Exception as e:
    print("bench $1 failed", e); print(open("$O/bench_$1.err").read()[-1500:])
PY
}
bench copies4
for v in 2 1 8; do
  ctas=2; if [ $v = 8 ]; then ctas=1; fi
  APOP_NVCC_DEFS="-DAPOP_NTAB_COPIES=$v -DAPOP_PROP_CTAS=$ctas" python anglican_b200/build.py --force > $O/build_$v.log 2>&1 || { echo build $v failed; tail -5 $O/build_$v.log; continue; }
  bench copies$v
done
python anglican_b200/build.py --force > /dev/null 2>&1
timeout 200 python tools/dbg_resolve.py > $O/dbg_resolve.txt 2>&1; head -6 $O/dbg_resolve.txt

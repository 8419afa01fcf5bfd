#!/bin/bash
# one-GPU job: full GPU suite, headline bench (short T), Box-Muller A/B, configs 2 and 5
R=${1:-r03e}
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
          d.get("kalman_check"), (d.get("roofline") or {}).get("frac"))
except Exception as e:
    print("bench $1 failed", e); print(open("$O/bench_$1.err").read()[-1500:])
PY
}
timeout 600 python bench.py --T 400 --steps 3 --warmup 3 > $O/bench_sv400.json 2> $O/bench_sv400.err; show sv400
APOP_NVCC_DEFS="-DAPOP_NORMAL_BM=1" python anglican_b200/build.py --force > $O/build_bm.log 2>&1
timeout 600 python bench.py --T 200 --steps 2 --warmup 3 --no-microbench --no-cpu-baseline > $O/bench_bm.json 2> $O/bench_bm.err; show bm
python anglican_b200/build.py --force > /dev/null 2>&1
timeout 600 python bench.py --workload lgssm1d --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err; show c2
APOP_GRAPH=0 timeout 600 python bench.py --workload lgssm1d --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c2_nograph.json 2> $O/bench_c2_nograph.err; show c2_nograph
timeout 600 python bench.py --workload lgmvn64 --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err; show c5

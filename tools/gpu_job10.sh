#!/bin/bash
# one-GPU: ncu time + DRAM bytes of every column-level kernel at 2^24 particles
R=${1:-r03k}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
python tools/prof_columns.py 24 > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file $O/columns.csv python tools/prof_columns.py 24 > $O/ncu.log 2>&1
tail -2 $O/plain.log; tail -2 $O/ncu.log
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("$O/columns.csv")) if len(r) > 8 and r[0].isdigit()]
# columns: ID, PID, proc, host, kernel, ctx, stream, block, grid, dev, cc, section, metric, unit, value
agg = collections.OrderedDict()
for r in rows:
    name = r[4].split("(")[0].replace("void ", "")[:44]
    agg.setdefault(name, {}).setdefault(r[-3], []).append((r[-2], float(r[-1].replace(",", ""))))
def val(d, key):
    v = d.get(key)
    if not v: return 0.0
    unit, x = v[-1]                      # the second (warm) pass
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "second": 1}.get(unit, 1)
print("%-46s %9s %9s %8s" % ("kernel (last launch)", "us", "MB dram", "GB/s"))
for k, d in agg.items():
    t = val(d, "gpu__time_duration.sum"); b = val(d, "dram__bytes_read.sum") + val(d, "dram__bytes_write.sum")
    print("%-46s %9.1f %9.1f %8.0f" % (k, t * 1e6, b / 1e6, b / t / 1e9 if t else 0))
PY

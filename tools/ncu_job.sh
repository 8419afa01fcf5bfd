#!/bin/bash
# ncu --set full of the three streaming kernels of the sweep at 2^27 particles (one GPU)
R=${1:-r03n}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
python tools/prof_sweep.py 27 6 > $O/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source off -k 'regex:k_propagate|k_fold_agg|k_ancestors$' -s 6 -c 6 \
    -o $O/prof python tools/prof_sweep.py 27 6 > $O/ncu.log 2>&1
tail -3 $O/plain.log; tail -3 $O/ncu.log
ncu -i $O/prof.ncu-rep --page raw --csv > $O/raw.csv 2> /dev/null
python tools/ncu_summary.py $O/raw.csv > $O/summary.csv; cat $O/summary.csv | cut -c1-400
ls -la $O

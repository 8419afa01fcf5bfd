#!/bin/bash
# one-GPU: chain phase timings and per-hard-tile cycles at three population sizes
R=${1:-r03i}
O=gpurun_out/$R
mkdir -p $O
python anglican_b200/build.py > /dev/null 2>&1
for lg in 27 24 20; do
  timeout 200 python tools/dbg_resolve.py $((1 << lg)) > $O/dbg_resolve_$lg.txt 2>&1; echo "== 2^$lg"; cat $O/dbg_resolve_$lg.txt
done

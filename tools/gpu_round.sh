#!/bin/bash
# First GPU call of the next round: things written after the GPU budget of round 1
# was spent.  One GPU:
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh one'
# Eight GPUs (each leg under its own timeout: more than 3 sweeps in flight hung
# once, profiles/README.md r02b):
#   gpurun --gpus 8 --timeout 1200 -- 'bash tools/gpu_round.sh eight'
R=${ROUND_TAG:-r03a}
O=gpurun_out/$R
mkdir -p $O
if [ "$1" = "one" ]; then
  # 1. whole GPU suite; the test_zz_* files hold what has never run on a GPU:
  #    dominant-weight populations (k_finalize prefix fix), IPMCMC node runner
  timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; tail -5 $O/pytest.log
  # 2. the shapes that used to fail with bits 0x204 / 0x604
  timeout 300 python tools/dbg_giant.py > $O/dbg_giant.txt 2>&1; cat $O/dbg_giant.txt
  # 3. the default bench line (2 sweeps in flight, T=2000) and the reference arm
  timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -c 600 $O/bench_default.json
  timeout 600 python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err
else
  N=${NGPU:-8}
  tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
  tr 600 29512 bench.py --gpus $N --T 400 --steps 2 --warmup 3 > $O/sv_${N}_f2.json 2> $O/sv_${N}_f2.err
  tr 600 29513 bench.py --gpus $N --T 400 --steps 2 --warmup 3 --inflight 3 > $O/sv_${N}_f3.json 2> $O/sv_${N}_f3.err
  # hypothesis: the stalls beyond 3 in flight come from engine streams sharing a hardware queue
  tr 300 29514 bench.py --gpus $N --T 400 --steps 2 --warmup 3 --inflight 4 --max-connections 32 > $O/sv_${N}_f4_mc32.json 2> $O/sv_${N}_f4_mc32.err
  # config 4 over the ranks (anglican_b200/ipmcmc.py)
  tr 600 29515 tools/bench_configs.py c4dist --nodes 32 --T 200 --steps 2 --warmup 1 > $O/c4dist_${N}.json 2> $O/c4dist_${N}.err
  for f in $O/*_${N}*.json; do echo $f; tail -c 400 $f; echo; done
fi

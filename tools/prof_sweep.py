"""Short SV sweep at full size for ncu: python tools/prof_sweep.py [log2 particles] [T]"""
import sys
sys.path.insert(0, '.')
import numpy as np
from anglican_b200 import Population
from anglican_b200.models import StochasticVolatility
from bench import sv_observations
n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 27)
T = int(sys.argv[2]) if len(sys.argv) > 2 else 6
y = sv_observations(2000)[:T]
with Population(n, seed=3) as pop:
    out = pop.sweep(StochasticVolatility(), y, want_final=False, want_stats=False, use_graph=-1)
    print("log_z", out.log_z)

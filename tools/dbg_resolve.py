import sys, ctypes as C, numpy as np
sys.path.insert(0, '.')
from anglican_b200 import Population, _lib as L
from anglican_b200.models import StochasticVolatility
from bench import sv_observations
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 27
T = 12
y = sv_observations(2000)[:T]
m = StochasticVolatility()
with Population(n, seed=3) as pop:
    for T_ in (4, 8, 12):
        pop.sweep(m, y[:T_], want_final=False, want_stats=False)
        out = np.zeros(8)
        pop.lib.apop_debug_resolve_phases(pop.ctx, out.ctypes.data_as(L.c_double_p))
        print("sweep T=%d: lap0 %.0f us, mid %.0f us, lap1 %.0f us, nseg %d, hard0 %d hard1 %d skip1 %d general folds %d" %
              (T_, out[0]/1e3, out[1]/1e3, out[2]/1e3, out[3], out[4], out[5], out[6], out[7]))
    lw = np.random.default_rng(7).normal(0, 3, n)
    col = pop.column(L.F64, data=lw)
    for u in (0.1, 0.9):
        anc, nl, j = pop.resample(col, u=u, order=L.ORDER_SORTED)
        out = np.zeros(8)
        pop.lib.apop_debug_resolve_phases(pop.ctx, out.ctypes.data_as(L.c_double_p))
        print("resample u=%.1f: lap0 %.0f us, mid %.0f us, lap1 %.0f us, nseg %d, hard0 %d hard1 %d skip1 %d general folds %d" %
              (u, out[0]/1e3, out[1]/1e3, out[2]/1e3, out[3], out[4], out[5], out[6], out[7]))

    tr = np.zeros(512)
    pop.lib.apop_debug_chain_trace(pop.ctx, tr.ctypes.data_as(L.c_double_p))
    nh = int(out[4])
    print("lap0 hard-tile fold cycles:", [int(tr[2*h]) for h in range(nh)])
    print("lap0 cycles between hard-tile starts:", [int(tr[2*h+3] - tr[2*h+1]) for h in range(nh-1)])
    print("general fold of tile 0 (cycles): weights %d, literal head %d, parallel fold %d in %d iterations" %
          (tr[480], tr[481], tr[482], tr[483]))

import sys; sys.path.insert(0,'.')
import numpy as np
from anglican_b200 import Population, _lib as L
from oracle import pyoracle as orc
rng = np.random.default_rng(1)
cases = []
for n in (450000, 600000, 800000, 1<<20):
    for frac in (0.6, 0.75, 0.95):
        cases.append((n, 1e-30, int(n*frac), 1.0, False))
cases.append((1<<20, 1e-30, 777777, 1.0, True))
cases.append((1<<20, 1e-5, 777777, 1e6, False))
cases.append((1<<20, 1e-30, 777777, 1.5, False))
cases.append((1<<22, 1e-30, (1<<22)-5, 3.0, False))
for n, small, pos, g, rnd in cases:
    w = np.full(n, small) * (rng.random(n) + 0.5 if rnd else 1.0); w[pos] = g
    with Population(n) as pop:
        wcol = pop.column(L.F64, data=w); anc = pop.column(L.I32)
        try:
            _, total, _ = pop.resample_weights(wcol, u=0.37, n_new=n, order=L.ORDER_EMISSION, anc=anc)
            ref = orc.resample_weights(w, n, 0.37)
            print(n, small, pos, g, rnd, pos//4096, 'ok', total == ref['total'], np.array_equal(anc.download(), ref['emission']))
        except Exception as e:
            print(n, small, pos, g, rnd, pos//4096, 'ERR', str(e)[:80])

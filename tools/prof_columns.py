"""Column-level kernels (K7 sample*/observe*, K5 gather, anglican.stat, normalise/resample)
once each at 2^24 particles, for `ncu --metrics gpu__time_duration.sum,dram__bytes...`.
python tools/prof_columns.py [log2 particles]"""
import sys
sys.path.insert(0, '.')
import numpy as np
from anglican_b200 import Population, _lib as L
n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 24)
rng = np.random.default_rng(1)
with Population(n, seed=3) as pop:
    x = pop.column(L.F64, data=rng.normal(0, 1, n))
    p = pop.column(L.F64, data=rng.uniform(0.1, 0.9, n))
    lw = pop.column(L.F64, fill=0.0)
    out = pop.column(L.F64)
    k = pop.column(L.I32)
    for rep in range(2):                                   # first pass warms the module / L2
        pop.sample_normal(0.0, 1.0, out); pop.observe_normal(x, 1.0, 0.3, lw)
        pop.sample_gamma(2.5, 1.5, out); pop.observe_gamma(2.5, 1.5, p, lw)
        pop.sample_beta(2.0, 3.0, out); pop.observe_beta(2.0, 3.0, p, lw)
        pop.sample_uniform(0.0, 2.0, out); pop.observe_uniform(0.0, 2.0, p, lw)
        pop.sample_discrete([1.0, 2.0, 3.0], k); pop.observe_discrete([1.0, 2.0, 3.0], k, lw)
        for fam, pars in (("exponential", (1.5,)), ("poisson", (3.0,)), ("student-t", (4.0,))):
            pop.sample_dist(fam, pars, out); pop.observe_dist(fam, pars, p, lw)
        m, lm, ess = pop.normalize(lw)
        lw.upload(rng.normal(0, 3, n))
        anc, nl, j = pop.resample(lw, u=0.37, order=L.ORDER_SORTED)
        pop.gather(anc, [x])
        pop.gather(anc, [k])
        pop.empirical_moments(x, lw)
        lw.fill(0.0)
    d = 8
    xv = pop.column(L.F64, width=d, data=rng.normal(0, 1, (n // 8, d)).repeat(8, axis=0)[:n])
    cov = 0.5 * np.eye(d) + 0.5 / d
    pop.observe_mvn(np.zeros(d), cov, xv, lw)
    pop.sample_mvn(np.zeros(d), cov, xv)
    pop.gather(anc, [xv])
    pop.synchronize()
    print("ok", m, ess)

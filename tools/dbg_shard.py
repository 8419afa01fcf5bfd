import sys, numpy as np
sys.path.insert(0, '.')
from anglican_b200 import Population, PopulationGroup
from anglican_b200.models import LinearGaussian1D, StochasticVolatility
rng = np.random.default_rng(20240601)
xs, ys = rng.normal(), []
for _ in range(30):
    xs = 0.9 * xs + rng.normal()
    ys.append(xs + rng.normal())
y = np.array(ys)
m = LinearGaussian1D()
for world, nl in [(2, 32768), (2, 4096), (2, 8)]:
    with Population(nl * world, seed=7) as pop:
        ref = pop.sweep(m, y, want_filt=False)
    with PopulationGroup(nl, world, seed=7) as grp:
        outs = grp.sweep(m, y)
    print(world, nl, ref.log_z, outs[0].log_z, outs[1].log_z)
    d = np.nonzero(ref.log_mean_w != outs[0].log_mean_w)[0]
    print(" first diverging step", d[:5], ref.log_mean_w[:4], outs[0].log_mean_w[:4])
    print(" ess", ref.ess[:4], outs[0].ess[:4])
    x = np.concatenate([o.x_final for o in outs])
    print(" x equal", np.array_equal(x, ref.x_final), np.sum(x != ref.x_final))

"""The whole sweep as one CUDA graph (apop_sweep_opts.use_graph, csrc/ssm.cu
run_graph): a replayed sweep must equal the kernel-by-kernel sweep bit for bit -
same particles, same log Z-hat, same lineages - also when the graph captured for
one sweep is replayed with another RNG key, other observations and another
retained path (the loop of `infer :smc` / :pgibbs, smc.cljc:58-63, pgibbs.cljc:117-126)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _models():
    from anglican_b200.models import GaussianHMM, LinearGaussian1D, StochasticVolatility
    init, trans = [1.0, 1.0, 1.0], [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]]
    return [("lgssm", LinearGaussian1D()), ("sv", StochasticVolatility()),
            ("hmm", GaussianHMM(init, trans, [-1.0, 1.0, 0.0], [1.0, 1.0, 1.0]))]


def _same(a, b, ctx):
    assert a.log_z == b.log_z, ctx
    assert np.array_equal(a.log_mean_w, b.log_mean_w) and np.array_equal(a.ess, b.ess), ctx
    assert np.array_equal(a.x_final, b.x_final), ctx
    if a.filt_mean is not None:
        assert np.array_equal(a.filt_mean, b.filt_mean), ctx


@pytest.mark.parametrize("n", [1000, 4096 * 3 + 5, 1 << 17])
def test_graph_replay_equals_direct_sweeps(n):
    from anglican_b200 import Population
    rng = np.random.default_rng(n)
    for name, m in _models():
        with Population(n, seed=1) as pg, Population(n, seed=1) as pd:
            for it in range(4):
                y = rng.normal(0, 1.0, 14)                 # new observations every sweep
                for p in (pg, pd):
                    p.set_seed(100 + it)
                a = pg.sweep(m, y, use_graph=1, want_filt=True)
                b = pd.sweep(m, y, use_graph=-1, want_filt=True)
                _same(a, b, (name, n, it))


@pytest.mark.parametrize("n", [64, 5000])
def test_graph_replay_of_conditional_sweeps_and_lineages(n):
    from anglican_b200 import Population
    rng = np.random.default_rng(7)
    for name, m in _models():
        y = rng.normal(0, 1.0, 12)
        T = y.size
        with Population(n, seed=2) as pg, Population(n, seed=2) as pd:
            retained = None
            for it in range(4):
                cond = retained is not None
                outs = []
                for p, g in ((pg, 1), (pd, -1)):
                    p.set_seed(300 + it)
                    outs.append(p.sweep(m, y, retained=retained, csmc=cond, n_cond_steps=T - 1,
                                        reset_weights=True, keep_history="compact", use_graph=g))
                _same(outs[0], outs[1], (name, n, it))
                k = (it * 37) % n
                ta, tb = outs[0].trace(k), outs[1].trace(k)
                assert np.array_equal(ta, tb), (name, n, it, "lineage")
                retained = ta                              # switches the shape once: recapture


def test_importance_sweep_in_a_graph():
    from anglican_b200 import Population
    from anglican_b200.models import LinearGaussian1D
    y = np.linspace(-1, 1, 9)
    with Population(3000, seed=5) as pg, Population(3000, seed=5) as pd:
        a = pg.sweep(LinearGaussian1D(), y, no_resample=True, use_graph=1)
        b = pd.sweep(LinearGaussian1D(), y, no_resample=True, use_graph=-1)
    assert a.log_z == b.log_z and np.array_equal(a.lw_final, b.lw_final)
    assert np.array_equal(a.x_final, b.x_final)

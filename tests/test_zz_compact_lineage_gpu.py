"""Compact lineages (keep_history = 2, csrc/ssm.cu `k_pack_lineage` /
`k_trace_compact`): two bits per particle and step instead of the 12 bytes of the
full history, the states of the ONE traced lineage (the `::past-samples` of
pgibbs.cljc:29-35, chosen by `rand-nth`, 122-123) recomputed from the counter-based
RNG.  Every traced path must equal the full-history trace bit for bit - plain,
conditional (pgibbs.cljc:76-110) and with ancestor sampling (pgas.cljc:107-131)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _models():
    from anglican_b200.models import GaussianHMM, LinearGaussian1D, StochasticVolatility
    rng = np.random.default_rng(5)
    init, trans = [1.0, 1.0, 1.0], [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]]
    return [("lgssm", LinearGaussian1D(), rng.normal(0, 1.5, 20)),
            ("sv", StochasticVolatility(), rng.normal(0, 0.7, 20)),
            ("hmm", GaussianHMM(init, trans, [-1.0, 1.0, 0.0], [1.0, 1.0, 1.0]),
             rng.normal(0, 1.2, 20))]


@pytest.mark.parametrize("n", [2, 6, 33, 4097, 50000])
@pytest.mark.parametrize("mode", ["smc", "csmc", "pgas"])
def test_compact_trace_equals_full_history_trace(n, mode):
    from anglican_b200 import Population
    for name, m, y in _models():
        T = y.size
        with Population(n, seed=9) as pop:
            retained = None
            for it in range(3):
                cond = mode != "smc" and retained is not None
                kw = dict(retained=retained if cond else None, csmc=cond,
                          n_cond_steps=T if mode == "pgas" else T - 1, reset_weights=mode != "smc",
                          ancestor_sampling=cond and mode == "pgas")
                slots = sorted({0, 1, n // 2, n - 1})
                pop.set_seed(700 + it)
                full = pop.sweep(m, y, keep_history=True, **kw)
                ref = [full.trace(k) for k in slots]
                pop.set_seed(700 + it)
                comp = pop.sweep(m, y, keep_history="compact", **kw)
                assert comp.log_z == full.log_z and np.array_equal(comp.x_final, full.x_final)
                for k, r in zip(slots, ref):
                    got = comp.trace(k)
                    assert np.array_equal(got, r), (name, mode, n, it, k)
                retained = ref[-1]


def test_compact_history_is_small_and_cannot_be_downloaded():
    from anglican_b200 import Population, _lib as L
    from anglican_b200.models import StochasticVolatility
    with Population(1 << 16, seed=1) as pop:
        out = pop.sweep(StochasticVolatility(), np.linspace(-1, 1, 8), keep_history="compact")
        with pytest.raises(L.ApopError):
            out.history()

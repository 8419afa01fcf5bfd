"""anglican_b200/ipmcmc.py on the GPU: nodes of one rank in flight on several
populations (context + stream each).  Not yet run on a GPU when written (the
round's GPU budget was spent); named test_zz_* so that it runs after the rest
of the GPU suite.  The rank protocol itself is covered on the CPU
(tests/test_ipmcmc_dist_cpu.py)."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_nodes_in_flight_equal_nodes_one_after_the_other():
    """node sweeps are keyed by (seed, iteration, node): running 4 of them at a
    time on 4 populations gives the chain of running them one by one"""
    from anglican_b200.ipmcmc import ipmcmc_iterations
    from anglican_b200.models import StochasticVolatility
    from bench import sv_observations
    m, y = StochasticVolatility(), sv_observations(40)
    runs = {}
    for width in (1, 4):
        gen = ipmcmc_iterations(m, y, 4096, 8, 4, seed=99, device=0, width=width,
                                distributed=False)
        runs[width] = [(s["log_zs"].copy(), s["csmc_indices"].copy(),
                        [r.results.copy() for r in s["runs"]]) for _, s in zip(range(3), gen)]
        gen.close()
    for a, b in zip(runs[1], runs[4]):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        for ra, rb in zip(a[2], b[2]):
            assert np.array_equal(ra, rb)


def test_hmm_posterior_table_nodes_in_flight():
    """algorithm_test.cljc:102-151, 265-301 through the multi-population driver"""
    from anglican_b200.ipmcmc import infer_ipmcmc
    from tests.test_sweep_gpu import _hmm
    m, _, y, h = _hmm()
    acc, wsum = np.zeros((17, 3)), 0.0
    it = infer_ipmcmc(m, y, 100, 10, seed=31337, device=0, paths=True, width=4, distributed=False)
    for i, s in enumerate(it):
        if i >= 10000:
            break
        w = math.exp(s.log_weight)
        acc[np.arange(17), np.asarray(s.result).astype(int)] += w
        wsum += w
    it.close()
    acc /= wsum
    rms = math.sqrt(((acc - np.array(h["posterior"])) ** 2).sum() / (3 * 16))
    assert rms < h["rms_threshold"], rms

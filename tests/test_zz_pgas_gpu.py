"""`infer :pgas` on the device: conditional sweeps with the
retained particle's ancestor redrawn at every observe (pgas.cljc:107-131,
`ancestor_sampling` of apop_smc_sweep) against the oracle.

Ancestry (integers) and the HMM's integer states must be bit-exact.  Continuous
states (lgssm, sv) are compared at the suite's 1e-9 relative: the device draws its
normals and exponentials with its own polynomial kernels (fastmath.cuh, <= 2 ulp from libm, which
the oracle uses), as every other Gaussian comparison in tests/ does.  The largest
distance seen is printed in ulps so that a real kernel bug (>> 2 ulp) stands out."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def _ulps(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.spacing(np.maximum(np.abs(b), 1.0))))     # same scale as _rel


def _same_states(name, got, ref, ctx):
    if name == "hmm":                                   # integer states: exact
        assert np.array_equal(got, ref), ctx
    else:
        r = _rel(got, ref)
        assert r < 1e-9, ctx + (r, _ulps(got, ref))


def _models():
    from anglican_b200.models import GaussianHMM, LinearGaussian1D, StochasticVolatility
    from oracle import pyoracle as orc
    rng = np.random.default_rng(3)
    lg = LinearGaussian1D()
    sv = StochasticVolatility()
    init, trans = [1.0, 1.0, 1.0], [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]]
    mu, sd = [-1.0, 1.0, 0.0], [1.0, 1.0, 1.0]
    return [
        ("lgssm", lg, orc.Model(orc.SSM_LGSSM1D, **lg.params), rng.normal(0, 1.5, 12)),
        ("sv", sv, orc.Model(orc.SSM_SV, m0=0.0, s0=sv.c.s0, a=sv.c.a, b=0.0, q=sv.c.q,
                             beta=sv.c.beta), rng.normal(0, 0.7, 12)),
        ("hmm", GaussianHMM(init, trans, mu, sd),
         orc.Model(orc.SSM_HMM, dim=3, hmm_init=init, hmm_trans=trans, hmm_mu=mu, hmm_sd=sd),
         np.array([0.9, 0.8, 0.7, 0.0, -0.025, -5.0, -2.0, -0.1, 0.0, 0.13, 0.45, 6.0])),
    ]


@pytest.mark.parametrize("n", [6, 7, 64, 5000, 20000])    # 6, 7: history rows not 16-byte aligned
def test_ancestor_sampling_sweep_equals_oracle(n):
    from anglican_b200 import Population
    from oracle import pyoracle as orc
    for name, m, om, y in _models():
        T = y.size
        with Population(n, seed=1) as pop:
            retained = None
            for it in range(3):
                cond = retained is not None
                pop.set_seed(500 + it)
                out = pop.sweep(m, y, retained=retained, csmc=cond, n_cond_steps=T,
                                reset_weights=True, keep_history=True, ancestor_sampling=cond)
                ref = orc.smc_sweep(om, y, n, seed=500 + it, retained=retained, csmc=cond,
                                    n_cond_steps=T, reset_weights=True, keep_history=True,
                                    ancestor_sampling=cond)
                anc, x = out.history()
                assert np.array_equal(anc, ref["anc_hist"]), (name, n, it)
                _same_states(name, x[:, :, 0], ref["x_hist"][:, :, 0], (name, n, it, "x_hist"))
                _same_states(name, out.x_final, ref["x_final"][:, 0], (name, n, it, "x_final"))
                assert abs(out.log_z - ref["log_z"]) <= 1e-9 * abs(ref["log_z"]), (name, n, it)
                if name != "hmm":
                    print("pgas %s n=%d it=%d: max |x - x_ref| = %.2f ulp"
                          % (name, n, it, _ulps(x[:, :, 0], ref["x_hist"][:, :, 0])))
                if cond:
                    assert np.any(anc[:, 0] != 0), (name, n)       # the draw did something
                slot = (7 * it + 3) % n
                tr = out.trace(slot)
                _same_states(name, tr, orc.trace_lineage(ref, slot)[:, 0], (name, n, it, "trace"))
                retained = tr


def test_pgas_needs_a_conditional_sweep():
    from anglican_b200 import Population, _lib as L
    from anglican_b200.models import LinearGaussian1D
    with Population(64, seed=1) as pop:
        with pytest.raises(L.ApopError):
            pop.sweep(LinearGaussian1D(), np.zeros(4), ancestor_sampling=True)


def test_infer_pgas_matches_the_smoother():
    """few particles, long series: only ancestor sampling keeps refreshing x_0"""
    from anglican_b200.inference import infer
    from anglican_b200.models import LinearGaussian1D
    from tests.test_pgas_oracle_cpu import PAR, lgssm_series, rts_smoother
    y = lgssm_series(40)
    ms, Ps = rts_smoother(y, **PAR)
    n, iters, burn = 6, 3000, 200
    acc, acc2, k = np.zeros(41), np.zeros(41), 0
    it = infer("pgas", LinearGaussian1D(**PAR), y, number_of_particles=n, paths=True, seed=99)
    for i, s in enumerate(it):
        if i >= iters * n:
            break
        if i >= burn * n:
            r = np.asarray(s.result)
            acc += r
            acc2 += r * r
            k += 1
        assert s.log_weight == 0.0 or not math.isfinite(s.log_weight)
    it.close()
    mean = acc / k
    var = acc2 / k - mean * mean
    assert np.abs(mean - ms).max() < 0.15
    assert np.all(np.abs(var / Ps - 1.0) < 0.25)


@pytest.mark.parametrize("n", [5, 4096, 30011, 1 << 20])
def test_one_pointer_resample_is_a_discrete_draw(n):
    """the draw behind the ancestor: apop_resample with n_new = 1 on any weights"""
    from anglican_b200 import Population, _lib as L
    from oracle import pyoracle as orc
    from tests.weights_cases import families
    rng = np.random.default_rng(12 + n)
    with Population(n, seed=1) as pop:
        wcol = pop.column(L.F64)
        anc = pop.column(L.I32)
        for name, w in families(n, rng).items():
            w = np.ascontiguousarray(w, dtype=np.float64)
            wcol.upload(w)
            for u in (0.0, 0.25, 0.6180339887, 1 - 2 ** -53):
                ref = orc.resample_weights(w, 1, u)
                _, total, _ = pop.resample_weights(wcol, u=u, n_new=1, order=L.ORDER_EMISSION,
                                                   anc=anc)
                assert total == ref["total"], (name, u)
                assert anc.download()[0] == ref["emission"][0], (name, u)

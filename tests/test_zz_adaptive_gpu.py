"""Adaptive (ESS-threshold) resampling of the fused sweep - SURVEY 8(f) row 4.  The
reference resamples at every global observe (smc.cljc:41-43; no ESS threshold exists
in src/), so this is an engine extension: checked against the oracle's restatement of
the same rule (ancestors/log Z/weights) and against exact truths (Kalman, HMM
forward).  threshold -> 0+ never resamples (= importance sampling,
importance.cljc:16-22), threshold = 1 resamples at (almost) every observe."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def _models():
    from anglican_b200.models import GaussianHMM, LinearGaussian1D, StochasticVolatility
    from oracle import pyoracle as orc
    rng = np.random.default_rng(21)
    lg, sv = LinearGaussian1D(), StochasticVolatility()
    init, trans = [1.0, 1.0, 1.0], [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]]
    mu, sd = [-1.0, 1.0, 0.0], [1.0, 1.0, 1.0]
    return [("lgssm", lg, orc.Model(orc.SSM_LGSSM1D, **lg.params), rng.normal(0, 1.5, 24)),
            ("sv", sv, orc.Model(orc.SSM_SV, m0=0.0, s0=sv.c.s0, a=sv.c.a, b=0.0, q=sv.c.q,
                                 beta=sv.c.beta), rng.normal(0, 0.7, 24)),
            ("hmm", GaussianHMM(init, trans, mu, sd),
             orc.Model(orc.SSM_HMM, dim=3, hmm_init=init, hmm_trans=trans, hmm_mu=mu, hmm_sd=sd),
             rng.normal(0, 1.2, 24))]


@pytest.mark.parametrize("n", [1000, 4096 * 5 + 3])
@pytest.mark.parametrize("thr", [0.3, 0.6, 0.9])
def test_adaptive_sweep_matches_oracle(n, thr):
    from anglican_b200 import Population
    from oracle import pyoracle as orc
    kept = resampled = 0
    for name, m, om, y in _models():
        for graph in (1, -1):
            with Population(n, seed=11) as pop:
                out = pop.sweep(m, y, ess_threshold=thr, keep_history=True, use_graph=graph)
                anc, x = out.history()
            ref = orc.smc_sweep(om, y, n, seed=11, ess_threshold=thr, keep_history=True)
            ctx = (name, n, thr, graph)
            skipped = ref["log_mean_w"] == 0.0
            kept += int(skipped.sum())
            resampled += int((~skipped).sum())
            assert np.array_equal(out.log_mean_w == 0.0, skipped), ctx
            assert np.array_equal(anc, ref["anc_hist"]), ctx
            assert _rel(out.log_mean_w, ref["log_mean_w"]) < 1e-9, ctx
            assert _rel(out.ess, ref["ess"]) < 1e-9, ctx
            assert abs(out.log_z - ref["log_z"]) <= 1e-9 * abs(ref["log_z"]), ctx
            assert _rel(out.lw_final, ref["lw_final"]) < 1e-9, ctx
            if name == "hmm":
                assert np.array_equal(out.x_final, ref["x_final"][:, 0]), ctx
            else:
                assert _rel(out.x_final, ref["x_final"][:, 0]) < 1e-9, ctx
    assert kept > 0 and resampled > 0                    # both branches were exercised


def test_adaptive_estimates_the_right_marginal():
    from anglican_b200 import Population
    from anglican_b200.models import LinearGaussian1D
    from oracle import pyoracle as orc
    rng = np.random.default_rng(20240601)
    xs, ys = rng.normal(), []
    for _ in range(50):
        xs = 0.9 * xs + rng.normal()
        ys.append(xs + rng.normal())
    y = np.array(ys)
    m = LinearGaussian1D()
    exact = orc.kalman1d(orc.Model(orc.SSM_LGSSM1D, **m.params), y)[0]
    with Population(1 << 18, seed=3) as pop:
        for thr in (0.25, 0.5, 1.0):
            out = pop.sweep(m, y, ess_threshold=thr)
            assert abs(out.log_z - exact) < 0.1, (thr, out.log_z, exact)
            # weighted final particles: self-normalised mean == plain mean after a resample
            w = np.exp(out.lw_final - out.lw_final.max())
            assert np.isfinite(w).all() and w.sum() > 0


def test_threshold_zero_is_the_reference_rule_and_bad_thresholds_are_refused():
    from anglican_b200 import Population, _lib as L
    from anglican_b200.models import StochasticVolatility
    y = np.linspace(-1, 1, 10)
    with Population(5000, seed=2) as pop:
        a = pop.sweep(StochasticVolatility(), y)
        b = pop.sweep(StochasticVolatility(), y, ess_threshold=0.0)
        assert a.log_z == b.log_z and np.array_equal(a.x_final, b.x_final)
        with pytest.raises(L.ApopError):
            pop.sweep(StochasticVolatility(), y, ess_threshold=1.5)


def test_alternating_thresholds_on_one_population_recapture_the_graph():
    """The ESS threshold is a by-value argument of the captured k_chain launch, and a threshold
    > 0 adds the final log-sum-exp node: it belongs to the graph key.  One Population, graph
    replay forced, thresholds alternating (0 included): every sweep must equal the oracle's
    and the never-graphed sweep's, not the sweep that was captured first."""
    from anglican_b200 import Population
    from oracle import pyoracle as orc
    name, m, om, y = _models()[0]
    n = 4096 * 3 + 17
    with Population(n, seed=11) as pop, Population(n, seed=11) as plain:
        for thr in (0.3, 0.0, 0.9, 0.3, 0.0, 0.6):
            out = pop.sweep(m, y, ess_threshold=thr, use_graph=1)
            eager = plain.sweep(m, y, ess_threshold=thr, use_graph=-1)
            ref = orc.smc_sweep(om, y, n, seed=11, ess_threshold=thr)
            assert np.array_equal(out.log_mean_w == 0.0, ref["log_mean_w"] == 0.0), thr
            assert abs(out.log_z - ref["log_z"]) <= 1e-9 * abs(ref["log_z"]), thr
            assert out.log_z == eager.log_z, thr
            assert np.array_equal(out.log_mean_w, eager.log_mean_w), thr
            assert np.array_equal(out.x_final, eager.x_final), thr
            assert _rel(out.x_final, ref["x_final"][:, 0]) < 1e-9, thr
            if thr > 0.0:
                assert np.array_equal(out.lw_final, eager.lw_final), thr
                assert _rel(out.lw_final, ref["lw_final"]) < 1e-9, thr

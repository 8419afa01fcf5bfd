"""Pins the oracle (oracle/anglican_oracle.c) against every golden vector the
reference's own tests hold for the hot path and exact truths."""
import json
import math
import os

import numpy as np
import pytest
from scipy import stats

from oracle import pyoracle as orc

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_goldens.json")))


def test_mvn_goldens():                       # runtime_test.cljc:41-46, tol 1e-4
    for c in G["mvn"]:
        assert abs(orc.mvn_logpdf(c["x"], c["mean"], c["cov"]) - c["logpdf"]) < 1e-4


def test_chi_squared_goldens():               # runtime_chi_squared_test.cljc:76-85, tol 1e-3
    for c in G["chi_squared"]:
        assert abs(orc.gamma_logpdf(c["x"], c["nu"] * 0.5, 0.5) - math.log(c["pdf"])) < 1e-3
        assert abs(orc.gamma_logpdf_direct(c["x"], c["nu"] * 0.5, 0.5) - math.log(c["pdf"])) < 1e-12


def test_categorical_goldens():               # runtime_test.cljc:18-24, exact `=`
    w = G["categorical"]["weights"]
    assert orc.discrete_logpmf(w, 0) == math.log(1.0 / 3.0)
    assert orc.discrete_logpmf(w, -1) == -math.inf
    assert orc.discrete_logpmf(w, 2) == -math.inf


def test_densities_vs_scipy():
    rng = np.random.default_rng(0)
    for _ in range(3000):
        k, r = rng.uniform(0.05, 50), rng.uniform(0.05, 20)
        x = rng.gamma(k, 1 / r)
        if x * r < 1e-3:
            continue          # commons' overflow branch cancels for tiny y: see DESIGN.md
        ref = stats.gamma.logpdf(x, k, scale=1 / r)
        assert abs(orc.gamma_logpdf(x, k, r) - ref) <= 1e-11 * max(1, abs(ref))
        a, b = rng.uniform(0.1, 20, 2)
        xb = min(max(rng.beta(a, b), 1e-300), 1 - 1e-16)
        ref = stats.beta.logpdf(xb, a, b)
        assert abs(orc.beta_logpdf(xb, a, b) - ref) <= 1e-11 * max(1, abs(ref))
        mu, sd, xn = rng.normal(), rng.uniform(0.1, 9), rng.normal(0, 5)
        assert abs(orc.normal_logpdf(xn, mu, sd) - stats.norm.logpdf(xn, mu, sd)) <= 1e-12 * 50
    alpha = np.array([0.7, 2.0, 3.5, 1.0])
    v = rng.dirichlet(alpha)
    assert abs(orc.dirichlet_logpdf(v, alpha) - stats.dirichlet.logpdf(v, alpha)) < 1e-11
    A = rng.normal(size=(6, 6))
    cov = A @ A.T + np.eye(6)
    x, mu = rng.normal(size=6), rng.normal(size=6)
    assert abs(orc.mvn_logpdf(x, mu, cov) - stats.multivariate_normal(mu, cov).logpdf(x)) < 1e-10
    assert orc.uniform_logpdf(0.5, 0.3, 1.2) == math.log(1 / 0.9)
    assert orc.uniform_logpdf(1.3, 0.3, 1.2) == -math.inf
    assert orc.beta_logpdf(1.2, 2, 2) == -math.inf and orc.gamma_logpdf(-1.0, 2, 2) == -math.inf


def test_resample12_literal_vs_closed_form():
    """hand-checked 12-particle wheel:
    a_ref[k] = s[N'-1-k], s[j] = searchsorted_right(C, wrap(x_j))"""
    c = G["resample12"]
    w = np.array(c["weights"])
    r = orc.resample_weights(w, 12, c["u"])
    assert list(r["anc_ref"]) == c["anc_ref"]
    C_ = np.cumsum(w)
    x = r["total"] * c["u"] + r["step"] * np.arange(12)
    x = np.where(x >= r["total"], x - r["total"], x)
    s = np.searchsorted(C_, x, side="right")
    assert list(s[::-1]) == c["anc_ref"]
    srt, j0 = orc.emission_to_sorted(r["emission"], r["lap"])
    assert np.all(np.diff(srt) >= 0) and j0 == int(np.sum(r["lap"] == 0))


def test_particle_weights_rules():            # smc.cljc:70-84
    w, m = orc.particle_weights([-1.0, -np.inf, 0.5, 0.5, -3.0])
    assert m == 0.5 and list(w[[1, 2, 3]]) == [0.0, 1.0, 1.0]
    assert w[0] == math.exp(-1.5) and w[4] == math.exp(-3.5)
    w, m = orc.particle_weights([-np.inf] * 4)          # all -inf: uniform
    assert m == -np.inf and list(w) == [1.0] * 4
    anc, new_lw = orc.resample([-np.inf] * 4, 4, 0.3)
    assert new_lw == -np.inf                            # smc.cljc:98-103
    with pytest.raises(ValueError):
        orc.particle_weights([0.0, np.nan])


def test_lse_log_marginal_ess_accept():
    assert orc.log_sum_exp(-np.inf, -np.inf) == -np.inf          # runtime.cljc:136
    assert abs(orc.log_sum_exp(1.0, 2.0) - np.logaddexp(1.0, 2.0)) < 1e-15
    rng = np.random.default_rng(2)
    lw = rng.normal(-20, 3, 1000)
    ref = np.log(np.mean(np.exp(lw - lw.max()))) + lw.max()
    assert abs(orc.log_marginal(lw) - ref) < 1e-12
    w = np.exp(lw - lw.max())
    assert abs(orc.ess(lw) - w.sum() ** 2 / (w ** 2).sum()) < 1e-9  # dist_test.cljc:83-94
    assert orc.accept(-5.0, -np.inf, 0.999)                         # lmh.cljc:146-151
    assert orc.accept(-1.0, -2.0, 0.5) and not orc.accept(-3.0, -2.0, 0.5)


def test_ipmcmc_gibbs_matches_reference_test():
    """ipmcmc_test.clj:33-45: frequencies within 0.005 with the intended rule;
    the literal `(> 0.0 max)` rule fails it."""
    c = G["ipmcmc"]
    lz, P, truth = np.array(c["log_zs"], float), c["P"], np.array(c["true_p_sample"])
    rng = np.random.default_rng(123)
    reps = 20000
    for as_written, should_pass in ((False, True), (True, False)):
        cnt, zs = np.zeros(lz.size), np.zeros(lz.size)
        for _ in range(reps):
            cs, z = orc.ipmcmc_gibbs(lz, P, rng.random(P), as_written=as_written)
            np.add.at(cnt, cs, 1)
            zs += z
        err = max(np.abs(cnt / cnt.sum() - truth).max(), np.abs(zs / (P * reps) - truth).max())
        assert (err < c["tolerance"]) == should_pass, (as_written, err)
    ps, lm = orc.norm_exp([-np.inf, -np.inf])
    assert list(ps) == [0.0, 0.0] and lm == 0.0                     # ipmcmc.cljc:92-95


def _hmm():
    h = G["hmm"]
    return orc.Model(orc.SSM_HMM, dim=3, hmm_init=h["init"], hmm_trans=h["trans"], hmm_mu=h["mu"],
                     hmm_sd=h["sd"]), np.array(h["observations"], float), h


def test_hmm_exact_truths():
    m, y, h = _hmm()
    assert abs(orc.hmm_forward_logz(m, y) - h["exact_log_z"]) < 1e-10
    assert np.abs(orc.hmm_smooth(m, y) - np.array(h["posterior"])).max() < 6e-5   # 4 dp table


def test_hmm_smc_posterior_table():
    """algorithm_test.cljc:102-151, 265-301: `:smc` 100 particles, 10 000 samples,
    RMS-L2 over the 17x3 table < 0.1 — here through the oracle's sweep."""
    m, y, h = _hmm()
    n, sweeps = 100, 100
    acc = np.zeros((17, 3))
    for s in range(sweeps):
        r = orc.smc_sweep(m, y, n, seed=1000 + s, keep_history=True)
        for k in range(n):
            tr = orc.trace_lineage(r, k)[:, 0].astype(int)
            acc[np.arange(17), tr] += 1
    acc /= n * sweeps
    rms = math.sqrt(((acc - np.array(h["posterior"])) ** 2).sum() / (3 * 16))
    assert rms < h["rms_threshold"], rms
    lz = [orc.smc_sweep(m, y, 1000, seed=s, want_final=False)["log_z"] for s in range(40)]
    assert abs(np.log(np.mean(np.exp(np.array(lz) - h["exact_log_z"])))) < 0.1   # unbiased Z-hat


def test_lgssm_smc_vs_kalman():
    rng = np.random.default_rng(20240601)
    T = 50
    x, ys = rng.normal(), []
    for _ in range(T):
        x = 0.9 * x + rng.normal()
        ys.append(x + rng.normal())
    m = orc.Model(orc.SSM_LGSSM1D, m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)
    exact, fm = orc.kalman1d(m, ys)
    r = orc.smc_sweep(m, ys, 20000, seed=5)
    assert abs(r["log_z"] - exact) < 0.1
    assert np.abs(r["filt_mean"][:, 0] - fm).max() < 0.05
    assert abs(r["log_z"] - r["log_mean_w"].sum()) < 1e-9


def test_csmc_keeps_retained_path():
    rng = np.random.default_rng(3)
    T = 12
    ys = rng.normal(size=T)
    m = orc.Model(orc.SSM_LGSSM1D, a=0.9, q=1.0, r=1.0)
    ret = rng.normal(size=T + 1)
    r = orc.smc_sweep(m, ys, 64, seed=9, retained=ret, csmc=True, n_cond_steps=T,
                      reset_weights=True, keep_history=True)
    assert np.allclose(orc.trace_lineage(r, 0)[:, 0], ret)        # slot 0 = retained (pgibbs.cljc:101-102)
    assert np.all(r["anc_hist"][:, 0] == 0)


def test_samplers_distributional():
    n = 4000
    g = np.array([orc.gamma_sample(2.1, 1.4, 7, s) for s in range(n)])
    assert stats.kstest(g, stats.gamma(2.1, scale=1 / 1.4).cdf).pvalue > 1e-3
    g = np.array([orc.gamma_sample(0.4, 2.0, 7, s) for s in range(n)])
    assert stats.kstest(g, stats.gamma(0.4, scale=0.5).cdf).pvalue > 1e-3
    for a, b in ((3.2, 2.2), (0.6, 0.8), (0.5, 3.0)):
        x = np.array([orc.beta_sample(a, b, 11, s) for s in range(n)])
        assert stats.kstest(x, stats.beta(a, b).cdf).pvalue > 1e-3, (a, b)
    z = np.array([orc.normal_at(3, s, 1, 1) for s in range(n)])
    assert stats.kstest(z, stats.norm.cdf).pvalue > 1e-3
    d = np.array([orc.dirichlet_sample([1.5, 0.7, 3.0], 5, s) for s in range(n)])
    assert np.abs(d.mean(0) - np.array([1.5, 0.7, 3.0]) / 5.2).max() < 0.02


def test_stat_goldens():
    """stat_test.cljc:16-21, 92-108: empirical-expectation / -mean of x = 0..10
    weighted by observe*(normal 5 2) is 5.0 +- 1e-12; the empirical distribution
    sums to 1 +- 1e-12."""
    from oracle import pyoracle as orc
    weighted = [(float(x), orc.normal_logpdf(float(x), 5.0, 2.0)) for x in range(11)]
    assert abs(orc.empirical_expectation(lambda v: v, weighted) - 5.0) < 1e-12
    assert abs(orc.empirical_mean(weighted) - 5.0) < 1e-12
    dist = orc.empirical_distribution(weighted)
    assert abs(sum(dist.values()) - 1.0) < 1e-12
    # symmetric weights: odd central moments vanish, variance below the prior's 4
    assert abs(float(orc.empirical_skew(weighted))) < 1e-12
    assert 3.0 < float(orc.empirical_variance(weighted)) < 4.0
    cb = orc.collect_by([0, 1, 0, 2], [0.0, 0.0, 0.0, 0.0])
    assert abs(cb[0] - math.log(0.5)) < 1e-15 and abs(cb[1] - math.log(0.25)) < 1e-15


def test_more_distributions_vs_scipy():
    """commons-math log-densities of the remaining scalar families
    (runtime.cljc:241-475, parameters of dist_test.cljc:334-396) vs scipy, 1e-12"""
    import scipy.stats as st
    from oracle import pyoracle as orc
    xs = [0.0, 0.3, 1.0, 2.5, 7.0]
    ks = [0, 1, 2, 5, 9]
    cases = [
        ("exponential", (3.1,), xs, lambda x: st.expon(scale=1 / 3.1).logpdf(x)),
        ("exponential", (1.4,), xs, lambda x: st.expon(scale=1 / 1.4).logpdf(x)),
        ("laplace", (2.1, 1.4), xs + [-3.0], lambda x: st.laplace(2.1, 1.4).logpdf(x)),
        ("poisson", (2.1,), ks, lambda k: st.poisson(2.1).logpmf(k)),
        ("poisson", (3.2,), ks + [40], lambda k: st.poisson(3.2).logpmf(k)),
        ("bernoulli", (0.7,), [0, 1], lambda k: st.bernoulli(0.7).logpmf(k)),
        ("student-t", (3.1,), xs + [-2.0], lambda x: st.t(3.1).logpdf(x)),
        ("student-t-loc-scale", (2.5, -10.0, 2.0), xs + [-12.0],
         lambda x: st.t(2.5, -10.0, 2.0).logpdf(x)),
        ("chi-squared", (5.0,), xs[1:], lambda x: st.chi2(5).logpdf(x)),
        ("uniform-discrete", (2, 5), [2, 3, 4], lambda k: st.randint(2, 5).logpmf(k)),
        ("binomial", (5, 0.7), [0, 1, 3, 5], lambda k: st.binom(5, 0.7).logpmf(k)),
    ]
    for name, params, vals, truth in cases:
        for v in vals:
            got, ref = orc.DIST_LOGP[name](float(v), *params), float(truth(v))
            assert abs(got - ref) <= 1e-12 * max(1.0, abs(ref)), (name, params, v, got, ref)
    # outside the support
    assert orc.exponential_logpdf(-1.0, 2.0) == -math.inf
    assert orc.poisson_logpmf(1.5, 2.0) == -math.inf and orc.poisson_logpmf(-1.0, 2.0) == -math.inf
    assert orc.bernoulli_logpmf(2.0, 0.3) == -math.inf
    assert orc.uniform_discrete_logpmf(5.0, 2, 5) == -math.inf       # max is exclusive
    assert orc.binomial_logpmf(6.0, 5, 0.7) == -math.inf


def test_wishart_and_mvt_logpdf_goldens():
    """wishart observe*: the two MATLAB values of runtime_wishart_test.cljc:161-176 (1e-3);
    both families against scipy (1e-10)."""
    from scipy import stats
    assert abs(orc.wishart_logpdf(np.eye(2), 3.5, [[0.7, 0], [0, 0.07]]) - (-5.3950)) < 1e-3
    assert abs(orc.wishart_logpdf([[3.1, -0.3], [-0.3, 11.0]], 7.5, [[3, 2], [2, 7]]) - (-12.5273)) < 1e-3
    rng = np.random.default_rng(8)
    for p in (2, 3, 5):
        A = rng.normal(size=(p, p))
        V = A @ A.T + p * np.eye(p)
        B = rng.normal(size=(p, p))
        X = B @ B.T + np.eye(p)
        n = p + 2.5
        ref = stats.wishart.logpdf(X, n, V)
        assert abs(orc.wishart_logpdf(X, n, V) - ref) <= 1e-10 * abs(ref)
        mu, y = rng.normal(size=p), rng.normal(size=p) * 2
        ref = stats.multivariate_t.logpdf(y, mu, V, 4.5)
        assert abs(orc.mvt_logpdf(y, mu, V, 4.5) - ref) <= 1e-10 * abs(ref)

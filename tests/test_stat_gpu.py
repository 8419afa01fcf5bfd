"""anglican.stat on the device (stats.cu) vs the literal restatement of
stat.cljc in the oracle and the reference's own goldens (stat_test.cljc:16-21,
92-108: tolerance 1e-12)."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-12          # stat_test.cljc:16


@pytest.fixture(scope="module")
def orc():
    from oracle import pyoracle
    return pyoracle


def _rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def test_reference_golden_normal_weighted_mean(orc):
    """stat_test.cljc:19-21, 92-102: x = 0..10 weighted by observe*(normal 5 2) has mean 5"""
    from anglican_b200 import Population, _lib as L
    xs = np.arange(11.0)
    with Population(11) as pop:
        x = pop.column(L.F64, data=xs)
        lw = pop.column(L.F64, fill=0.0)
        pop.observe_normal(5.0, 2.0, x, lw)              # log-weights made on the device too
        mom = pop.empirical_moments(x, lw)
        _, pr = pop.collect_discrete(x, 11, lw)
    assert abs(mom["mean"][0] - 5.0) < TOL
    assert abs(pr.sum() - 1.0) < TOL                     # stat_test.cljc:104-108
    weighted = [(v, orc.normal_logpdf(v, 5.0, 2.0)) for v in xs]
    assert abs(orc.empirical_mean(weighted) - 5.0) < TOL
    assert abs(mom["variance"][0] - orc.empirical_variance(weighted)) < TOL


@pytest.mark.parametrize("n,width", [(1, 1), (1000, 1), (4097, 3), (20000, 11)])
def test_moments_match_oracle(orc, n, width):
    from anglican_b200 import Population, _lib as L
    rng = np.random.default_rng(n + width)
    xs = rng.normal(1.5, 2.0, (n, width)) * np.arange(1, width + 1)
    lws = rng.normal(0, 2.0, n)
    lws[rng.integers(0, n)] = -np.inf if n > 1 else 0.0  # a dead particle weighs nothing
    with Population(n) as pop:
        x = pop.column(L.F64, width=width, data=xs)
        lw = pop.column(L.F64, data=lws)
        mom = pop.empirical_moments(x, lw)
        mom_eq = pop.empirical_moments(x)                 # equal weights
    weighted = [(xs[i], lws[i]) for i in range(n)]
    assert _rel(mom["mean"], orc.empirical_mean(weighted)) < TOL
    assert _rel(mom_eq["mean"], xs.mean(0)) < TOL
    if n > 1:
        assert _rel(mom["variance"], orc.empirical_variance(weighted)) < 1e-11
        assert _rel(mom["std"], orc.empirical_std(weighted)) < 1e-11
        assert _rel(mom["skew"], orc.empirical_skew(weighted)) < 1e-10
        assert _rel(mom["kurtosis"], orc.empirical_kurtosis(weighted)) < 1e-10
        assert _rel(mom_eq["variance"], xs.var(0)) < 1e-11


@pytest.mark.parametrize("n,d", [(500, 2), (3000, 8), (5000, 64)])
def test_covariance_matches_oracle(orc, n, d):
    from anglican_b200 import Population, _lib as L
    rng = np.random.default_rng(d)
    A = rng.normal(size=(d, d)) / math.sqrt(d)
    xs = rng.normal(size=(n, d)) @ A + rng.normal(size=d)
    lws = rng.normal(0, 1.0, n)
    with Population(n) as pop:
        x = pop.column(L.F64, width=d, data=xs)
        lw = pop.column(L.F64, data=lws)
        mean, cov = pop.empirical_covariance(x, lw)
    w = np.exp(lws - lws.max())
    mu = (w[:, None] * xs).sum(0) / w.sum()
    ref = np.einsum("i,ij,ik->jk", w, xs - mu, xs - mu) / w.sum()
    assert _rel(mean, mu) < TOL
    assert _rel(cov, ref) < 1e-11
    if n <= 500:
        weighted = [(xs[i], lws[i]) for i in range(n)]
        assert _rel(cov, orc.empirical_covariance(weighted)) < 1e-11


def test_collect_discrete_matches_oracle(orc):
    from anglican_b200 import Population, _lib as L
    rng = np.random.default_rng(3)
    n, K = 7000, 5
    vals = rng.integers(0, K - 1, n)                     # value K-1 never occurs
    lws = rng.normal(0, 1.5, n)
    with Population(n) as pop:
        vi = pop.column(L.I32, data=vals.astype(np.int32))
        vf = pop.column(L.F64, data=vals.astype(np.float64))
        lw = pop.column(L.F64, data=lws)
        lw_i, pr_i = pop.collect_discrete(vi, K, lw)
        lw_f, pr_f = pop.collect_discrete(vf, K, lw)
        lw_eq, pr_eq = pop.collect_discrete(vi, K)
    ref = orc.collect_by(list(vals), list(lws))
    dist = orc.empirical_distribution(ref.items())
    for k in range(K - 1):
        assert abs(lw_i[k] - ref[k]) < 1e-11 and abs(lw_f[k] - ref[k]) < 1e-11
        assert abs(pr_i[k] - dist[k]) < 1e-12
        assert abs(pr_eq[k] - np.mean(vals == k)) < 1e-12
    assert lw_i[K - 1] == -np.inf and pr_i[K - 1] == 0.0
    assert abs(pr_i.sum() - 1.0) < TOL


def test_hmm_posterior_from_device_histogram(orc):
    """final-state posterior of BASELINE config 1 straight from the sweep's
    device column (no particle download) == histogram of the downloaded states"""
    from anglican_b200 import Population
    from anglican_b200.models import GaussianHMM
    y = [0.9, 0.8, 0.7, 0.0, -0.025, -5.0, -2.0, -0.1, 0.0, 0.13, 0.45, 6, 0.2, 0.3, -1, -1]
    m = GaussianHMM([1.0, 1.0, 1.0], [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]],
                    [-1, 1, 0], [1, 1, 1])
    with Population(50000, seed=12) as pop:
        out = pop.sweep(m, y)
        _, pr = pop.collect_discrete("final", 3)
        mom = pop.empirical_moments("final")
    host = np.array([(out.x_final == k).mean() for k in range(3)])
    assert np.max(np.abs(pr - host)) < 1e-12
    assert abs(mom["mean"][0] - out.x_final.mean()) < 1e-12
    # algorithm_test.cljc:131-147, last row of the exact table: [0.2..]-ish posterior; smoothing
    # truth for the final step equals the filtering truth
    truth = orc.hmm_smooth(orc.Model(orc.SSM_HMM, dim=3, hmm_init=[1, 1, 1],
                                     hmm_trans=[[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]],
                                     hmm_mu=[-1, 1, 0], hmm_sd=[1, 1, 1]), y)[-1]
    assert np.max(np.abs(pr - truth)) < 0.02

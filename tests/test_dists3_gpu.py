"""multivariate-t and wishart on the device (csrc/dists3.cu; runtime.cljc:371-401,
479-545): log-densities against the oracle's literal restatement (1e-9 relative),
which is pinned on runtime_wishart_test.cljc:161-176 and scipy in
tests/test_oracle_golden.py; samplers by the reference's own moment tests
(runtime_test.cljc:116-130, runtime_wishart_test.cljc:178-230)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
N = 100000


@pytest.fixture(scope="module")
def pop():
    from anglican_b200 import Population
    with Population(N, seed=161803) as p:
        yield p


def _rel(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


@pytest.mark.parametrize("d", [2, 3, 7])
def test_mvt_observe_matches_oracle(pop, d):
    from anglican_b200 import _lib as L
    from oracle import pyoracle as orc
    rng = np.random.default_rng(d)
    A = rng.normal(size=(d, d))
    sigma = A @ A.T + d * np.eye(d)
    mu = rng.normal(size=d)
    nu = 4.5
    vals = rng.normal(size=(N, d)) * 3
    x = pop.column(L.F64, width=d, data=vals)
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_mvt(nu, mu, sigma, x, lw)
    got = lw.download()
    ref = np.array([orc.mvt_logpdf(vals[k], mu, sigma, nu) for k in range(2000)])
    assert _rel(got[:2000], ref) < 1e-9
    # per-particle mean column, shared value
    lw.fill(0.0)
    pop.observe_mvt(nu, x, sigma, mu, lw)
    assert _rel(lw.download()[:2000], ref) < 1e-9


def test_mvt_sampler_moments(pop):
    """runtime_test.cljc:116-130: mean within 20 %, covariance * (nu-2)/nu within 50 % of sigma"""
    from anglican_b200 import _lib as L
    nu, mu = 5.0, np.array([3.0, 1.0])
    sigma = np.array([[1.0, 0.1], [0.1, 1.0]])
    x = pop.column(L.F64, width=2)
    pop.sample_mvt(nu, mu, sigma, x, draw=3)
    s = x.download().reshape(N, 2)
    assert np.all(np.abs(s.mean(0) - mu) <= 0.2 * mu)
    cov = np.cov(s.T) * (nu - 2) / nu
    assert np.all(np.abs(cov - sigma) <= 0.5 * sigma)
    # sharper: the CLT band of this sample size
    assert np.all(np.abs(s.mean(0) - mu) < 5 * np.sqrt(nu / (nu - 2) / N))


@pytest.mark.parametrize("n_df,V", [(3.5, [[0.7, 0.0], [0.0, 0.07]]), (7.5, [[3.0, 2.0], [2.0, 7.0]]),
                                    (9.0, np.eye(4) + 0.3)])
def test_wishart_observe_matches_oracle(pop, n_df, V):
    from anglican_b200 import _lib as L
    from oracle import pyoracle as orc
    V = np.asarray(V, float)
    p = V.shape[0]
    rng = np.random.default_rng(p)
    B = rng.normal(size=(N, p, p))
    vals = B @ np.transpose(B, (0, 2, 1)) + 0.5 * np.eye(p)
    vals[0] = np.eye(p)
    x = pop.column(L.F64, width=p * p, data=vals.reshape(N, p * p))
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_wishart(n_df, V, x, lw)
    got = lw.download()
    ref = np.array([orc.wishart_logpdf(vals[k], n_df, V) for k in range(2000)])
    assert _rel(got[:2000], ref) < 1e-9
    if p == 2 and n_df == 3.5:
        assert abs(got[0] - (-5.3950)) < 1e-3                  # runtime_wishart_test.cljc:165-169


@pytest.mark.parametrize("n_df,V,tm,tv", [
    (10, [[0.7, 0.0], [0.0, 0.07]], 0.3, 0.5), (10, [[3.0, 2.0], [2.0, 7.0]], 3.0, 100.0)])
def test_wishart_sampler_moments(pop, n_df, V, tm, tv):
    """runtime_wishart_test.cljc:178-230: mean n V, variance n (V_ij^2 + V_ii V_jj), by the
    reference's hand-picked thresholds and by the CLT band"""
    from anglican_b200 import _lib as L
    V = np.asarray(V, float)
    p = V.shape[0]
    x = pop.column(L.F64, width=p * p)
    pop.sample_wishart(n_df, V, x, draw=1)
    s = x.download().reshape(N, p, p)
    mean_th = n_df * V
    var_th = n_df * (V ** 2 + np.outer(np.diag(V), np.diag(V)))
    assert np.max(np.abs(s.mean(0) - mean_th)) < tm
    assert np.max(np.abs(s.var(0, ddof=1) - var_th)) < tv
    assert np.all(np.abs(s.mean(0) - mean_th) < 5 * np.sqrt(var_th / N) + 1e-12)
    assert np.allclose(s, np.transpose(s, (0, 2, 1)))          # symmetric, positive definite
    assert np.all(np.linalg.eigvalsh(s[:2000]) > 0)


def test_degenerate_wishart_df_and_aliasing_draw_ids_are_refused(pop):
    """n <= p - 1 puts a pole / negative argument into the multivariate gamma (runtime.cljc:479-485):
    observe refuses it like sample does.  draw ids >= 256 of the matrix families would reuse the
    counters of draw - 256: refused instead of silently repeating the numbers."""
    from anglican_b200 import _lib as L
    V = np.eye(3)
    x = pop.column(L.F64, width=9, data=np.tile(np.eye(3).reshape(9), (N, 1)))
    lw = pop.column(L.F64, fill=0.0)
    for bad in (2.0, 1.0, -1.0):
        with pytest.raises(L.ApopError) as e:
            pop.observe_wishart(bad, V, x, lw)
        assert e.value.code == -1 and "n > p - 1" in str(e.value)
    pop.observe_wishart(2.5, V, x, lw)                       # n > p - 1 is fine
    assert np.isfinite(lw.download()).all()
    v3 = pop.column(L.F64, width=3)
    for call in (lambda d: pop.sample_wishart(5.0, V, x, draw=d),
                 lambda d: pop.sample_mvt(4.0, np.zeros(3), V, v3, draw=d),
                 lambda d: pop.sample_mvn(np.zeros(3), V, v3, draw=d)):
        call(255)
        with pytest.raises(L.ApopError) as e:
            call(256)
        assert e.value.code == -1 and "draw_id" in str(e.value)

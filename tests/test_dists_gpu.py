"""observe*/sample* parity of the runtime distributions (runtime.cljc) through
the C ABI.  Log-densities: 1e-9 relative vs the oracle (north_star tolerance).
Samplers: distributional (moments / KS), since the reference RNG is unseeded
(runtime.cljc:142-147)."""
import json
import os

import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
N = 20000


@pytest.fixture(scope="module")
def orc():
    from oracle import pyoracle
    return pyoracle


@pytest.fixture(scope="module")
def pop():
    from anglican_b200 import Population
    p = Population(N, seed=0xA11CE)
    yield p
    p.close()


def _close(a, b, tol=1e-9):
    a, b = np.asarray(a), np.asarray(b)
    fin = np.isfinite(b)
    assert np.array_equal(np.isfinite(a), fin)
    assert np.array_equal(a[~fin], b[~fin]) or np.all(np.isnan(a[~fin]) == np.isnan(b[~fin]))
    err = np.abs(a[fin] - b[fin]) / np.maximum(1.0, np.abs(b[fin]))
    assert err.max(initial=0.0) <= tol, err.max()


def test_observe_normal(pop, orc):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(1)
    mu, sd, x = rng.normal(size=N), rng.uniform(0.1, 5, N), rng.normal(0, 4, N)
    base = rng.normal(size=N)
    lw = pop.column(L.F64, data=base)
    pop.observe_normal(pop.column(L.F64, data=mu), pop.column(L.F64, data=sd),
                       pop.column(L.F64, data=x), lw)
    ref = base + np.array([orc.normal_logpdf(a, b, c) for a, b, c in zip(x, mu, sd)])
    _close(lw.download(), ref)
    lw2 = pop.column(L.F64, fill=0.0)
    pop.observe_normal(5.0, 2.0, pop.column(L.F64, data=x), lw2)     # stat_test.cljc:19-21
    _close(lw2.download(), [orc.normal_logpdf(a, 5.0, 2.0) for a in x])


def test_observe_gamma_beta_uniform(pop, orc):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(2)
    k, r = rng.uniform(0.2, 30, N), rng.uniform(0.1, 10, N)
    x = rng.gamma(k, 1 / r)
    x[:10] = -1.0
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_gamma(pop.column(L.F64, data=k), pop.column(L.F64, data=r),
                      pop.column(L.F64, data=x), lw)
    got = lw.download()
    _close(got, [orc.gamma_logpdf_direct(a, b, c) for a, b, c in zip(x, k, r)], 1e-11)
    # commons-math's Lanczos-prefactor form agrees wherever it does not cancel (y >= 1e-3)
    sel = (x * r >= 1e-3)
    _close(got[sel], np.array([orc.gamma_logpdf(a, b, c) for a, b, c in zip(x, k, r)])[sel], 1e-9)

    a, b = rng.uniform(0.3, 20, N), rng.uniform(0.3, 20, N)
    xb = rng.beta(a, b)
    xb[:5] = [-0.1, 1.1, 0.5, 0.25, 0.75]
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_beta(pop.column(L.F64, data=a), pop.column(L.F64, data=b),
                     pop.column(L.F64, data=xb), lw)
    _close(lw.download(), [orc.beta_logpdf(p, q, s) for p, q, s in zip(xb, a, b)])

    lo, hi = rng.uniform(-2, 0, N), rng.uniform(0.5, 3, N)
    xu = rng.uniform(-3, 4, N)
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_uniform(pop.column(L.F64, data=lo), pop.column(L.F64, data=hi),
                        pop.column(L.F64, data=xu), lw)
    _close(lw.download(), [orc.uniform_logpdf(p, q, s) for p, q, s in zip(xu, lo, hi)])


def test_golden_densities(pop):
    """the reference's own golden vectors (tests/golden/reference_goldens.json)"""
    from anglican_b200 import _lib as L
    g = json.load(open(os.path.join(GOLD, "reference_goldens.json")))
    for case in g["mvn"]:
        lw = pop.column(L.F64, fill=0.0)
        pop.observe_mvn(case["mean"], case["cov"], case["x"], lw)
        assert abs(lw.download()[0] - case["logpdf"]) < 1e-4        # runtime_test.cljc:41-46
    for case in g["chi_squared"]:
        lw = pop.column(L.F64, fill=0.0)
        pop.observe_gamma(case["nu"] * 0.5, 0.5, case["x"], lw)    # runtime.cljc:333-340
        assert abs(lw.download()[0] - np.log(case["pdf"])) < 1e-3  # runtime_chi_squared_test.cljc:76-85
    c = g["categorical"]
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_discrete(c["weights"], 0, lw)
    assert lw.download()[0] == np.log(1.0 / 3.0)                   # runtime_test.cljc:18-24
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_discrete(c["weights"], -1, lw)
    assert lw.download()[0] == -np.inf


def test_discrete_dirichlet_mvn_observe(pop, orc):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(3)
    K = 5
    W = rng.uniform(0.1, 2, (N, K))
    v = rng.integers(-1, K + 1, N).astype(np.int32)
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_discrete(pop.column(L.F64, width=K, data=W), pop.column(L.I32, data=v), lw)
    with np.errstate(divide="ignore"):
        _close(lw.download(), [orc.discrete_logpmf(W[i], v[i]) for i in range(N)])
    alpha = rng.uniform(0.3, 4, K)
    V = rng.dirichlet(alpha, N)
    lw = pop.column(L.F64, fill=0.0)
    pop.observe_dirichlet(alpha, pop.column(L.F64, width=K, data=V), lw)
    _close(lw.download(), [orc.dirichlet_logpdf(V[i], alpha) for i in range(N)])
    for d in (2, 7, 64):
        A = rng.normal(size=(d, d))
        cov = A @ A.T / d + 0.5 * np.eye(d)
        X = rng.normal(size=(N, d))
        mean = rng.normal(size=d)
        lw = pop.column(L.F64, fill=0.0)
        pop.observe_mvn(mean, cov, pop.column(L.F64, width=d, data=X), lw)
        got = lw.download()
        ref = [orc.mvn_logpdf(X[i], mean, cov) for i in range(0, N, 97)]
        _close(got[::97], ref)
        _close(got[::97], stats.multivariate_normal(mean, cov).logpdf(X[::97]), 1e-9)


def test_samplers_distributional(pop):
    from anglican_b200 import _lib as L
    out = pop.column(L.F64)
    checks = []
    pop.sample_normal(0.3, 0.1, out, draw=0)                        # dist_test.cljc test-dists
    checks.append(("normal", out.download(), stats.norm(0.3, 0.1)))
    pop.sample_gamma(2.1, 1.4, out, draw=1)
    checks.append(("gamma", out.download(), stats.gamma(2.1, scale=1 / 1.4)))
    pop.sample_gamma(0.4, 2.0, out, draw=2)
    checks.append(("gamma<1", out.download(), stats.gamma(0.4, scale=0.5)))
    pop.sample_beta(3.2, 2.2, out, draw=3)
    checks.append(("beta", out.download(), stats.beta(3.2, 2.2)))
    pop.sample_uniform(0.3, 1.2, out, draw=4)
    checks.append(("uniform", out.download(), stats.uniform(0.3, 0.9)))
    for name, x, dist in checks:
        p = stats.kstest(x, dist.cdf).pvalue
        assert p > 1e-4, (name, p)
    iout = pop.column(L.I32)
    pop.sample_discrete([1.0, 3.0, 2.0], iout, draw=5)
    freq = np.bincount(iout.download(), minlength=3) / N
    assert np.abs(freq - np.array([1, 3, 2]) / 6).max() < 0.015
    dout = pop.column(L.F64, width=3)
    alpha = np.array([1.5, 0.7, 3.0])
    pop.sample_dirichlet(alpha, dout, draw=6)
    D = dout.download()
    assert np.allclose(D.sum(1), 1.0, atol=1e-12)
    assert np.abs(D.mean(0) - alpha / alpha.sum()).max() < 0.01
    mout = pop.column(L.F64, width=2)
    cov = np.array([[1.0, 0.5], [0.5, 1.0]])
    pop.sample_mvn([0.5, -0.2], cov, mout, draw=7)
    M = mout.download()
    assert np.abs(M.mean(0) - [0.5, -0.2]).max() < 0.03
    assert np.abs(np.cov(M.T) - cov).max() < 0.05


def test_fastmath_accuracy(pop):
    """csrc/fastmath.cuh (constant-bank polynomial exp/log/sqrt/sincospi/rcp) against
    numpy: <= 2 ulp (sin/cos: 2 ulp of 1.0 absolute, as for any reduced-argument form)."""
    from anglican_b200 import _lib as L
    import ctypes as C
    n = pop.n
    rng = np.random.default_rng(99)
    cin, cout = pop.column(L.F64), pop.column(L.F64)

    def run(fn, x):
        cin.upload(x)
        L.check(pop.ctx, pop.lib.apop_selftest_math(pop.ctx, fn, cin.handle, cout.handle))
        return cout.download()

    def ulps(a, b):
        return np.abs(a - b) / np.spacing(np.abs(b))

    x = np.concatenate([rng.uniform(-707, 707, n // 2), rng.normal(0, 3, n - n // 2)])
    x[:4] = [0.0, -0.0, -745.2, 709.0]                     # library fallback range too
    assert ulps(run(0, x), np.exp(x)).max() <= 2.0
    u = np.concatenate([rng.random(n // 2), 2.0 ** rng.uniform(-54, 0, n - n // 2)])
    u[u <= 0] = 0.5
    assert ulps(run(1, u), np.log(u)).max() <= 2.0
    a = 2.0 ** rng.uniform(-60, 8, n) * rng.uniform(1, 2, n)
    assert ulps(run(2, a), np.sqrt(a)).max() <= 1.0
    assert ulps(run(5, a), 1.0 / a).max() <= 1.0
    t = rng.uniform(0, 2, n)
    t[:5] = [0.0, 0.5, 1.0, 1.5, 2.0 - 2.0 ** -52]
    import mpmath as mp
    mp.mp.dps = 40
    sn, cs = run(3, t), run(4, t)
    for k in list(range(5)) + list(rng.integers(0, n, 3000)):
        tk = mp.mpf(float(t[k]))
        assert abs(mp.mpf(float(sn[k])) - mp.sinpi(tk)) <= 2.3e-16, (k, t[k])
        assert abs(mp.mpf(float(cs[k])) - mp.cospi(tk)) <= 2.3e-16, (k, t[k])

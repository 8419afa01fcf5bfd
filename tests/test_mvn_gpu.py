"""mvn on the fp64 tensor pipe (csrc/mvn.cu): the batched whitened log-density
(runtime.cljc:349-368) and the fused d-dimensional linear-Gaussian sweep of
BASELINE config 5, against the oracle (1e-9 relative, north_star) and the exact
Kalman filter."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _spd(d, rng):
    return 0.5 * np.eye(d) + 0.5 * np.ones((d, d)) / d + 0.05 * np.diag(rng.random(d))


@pytest.mark.parametrize("d", [8, 16, 32, 64])
def test_observe_mvn_tensor_path_matches_oracle(d):
    from anglican_b200 import Population, _lib as L
    from oracle import pyoracle as orc
    rng = np.random.default_rng(d)
    n = 1003                                   # ragged: not a multiple of 8
    cov = _spd(d, rng)
    mean = rng.normal(size=d)
    X = rng.normal(size=(n, d))
    M = rng.normal(size=(n, d)) * 0.1
    with Population(n, seed=1) as pop:
        lw = pop.column(L.F64, fill=0.25)
        pop.observe_mvn(mean, cov, pop.column(L.F64, width=d, data=X), lw)      # shared mean
        got = lw.download()
        lw2 = pop.column(L.F64, fill=0.0)
        pop.observe_mvn(pop.column(L.F64, width=d, data=M), cov, X[7], lw2)     # per-particle mean
        got2 = lw2.download()
    for i in range(0, n, 53):
        ref = orc.mvn_logpdf(X[i], mean, cov)
        assert abs(got[i] - (0.25 + ref)) <= 1e-9 * max(1.0, abs(ref))
        ref2 = orc.mvn_logpdf(X[7], M[i], cov)
        assert abs(got2[i] - ref2) <= 1e-9 * max(1.0, abs(ref2))


def _lgmvn(d, T, seed=20240603):
    rng = np.random.default_rng(seed)
    Q, _ = np.linalg.qr(rng.normal(size=(d, d)))
    A = 0.9 * Q
    Qc = 0.1 * np.eye(d)
    R = 0.5 * np.eye(d) + 0.5 * np.ones((d, d)) / d
    x = rng.normal(size=d)
    ys = []
    LR = np.linalg.cholesky(R)
    for _ in range(T):
        x = A @ x + np.sqrt(0.1) * rng.normal(size=d)
        ys.append(x + LR @ rng.normal(size=d))
    return A, Qc, R, np.array(ys)


@pytest.mark.parametrize("d,n", [(8, 1000), (64, 4096 + 40)])
def test_lgmvn_sweep_matches_oracle(d, n):
    from anglican_b200 import Population
    from anglican_b200.models import LinearGaussianMVN
    from oracle import pyoracle as orc
    T = 6
    A, Qc, R, y = _lgmvn(d, T)
    m = LinearGaussianMVN(A, Qc, R)
    om = orc.Model(orc.SSM_LGMVN, dim=d, s0=1.0, A=A, Lq=np.linalg.cholesky(Qc), R=R)
    with Population(n, seed=11) as pop:
        out = pop.sweep(m, y.ravel(), keep_history=True)
        anc, xh = out.history()
    ref = orc.smc_sweep(om, y.ravel(), n, seed=11, keep_history=True)
    assert np.max(np.abs(out.log_mean_w - ref["log_mean_w"]) /
                  np.maximum(1.0, np.abs(ref["log_mean_w"]))) < 1e-9
    assert abs(out.log_z - ref["log_z"]) <= 1e-9 * abs(ref["log_z"])
    # The log-weights agree to ~1e-12, not bit for bit (tensor-pipe summation order), so a
    # resampling pointer that falls within rounding distance of a cumulative edge may pick the
    # NEIGHBOURING ancestor.  Such boundary ties are counted, not averaged away: walk the
    # ancestry step by step; up to the first step with a tie every particle must agree with the
    # oracle (all of them, 1e-9), and at that step every differing slot must be a shift by
    # exactly one ancestor, at most one particle in 10^5.  After a tie the two runs are
    # different populations and are no longer compared particle by particle.
    ties, tie_step = 0, None
    for t in range(T):
        assert np.allclose(xh[t], ref["x_hist"][t], rtol=1e-9, atol=1e-9), (d, n, t)
        diff = anc[t] != ref["anc_hist"][t]
        if diff.any():
            ties, tie_step = int(diff.sum()), t
            shift = np.abs(anc[t][diff].astype(np.int64) - ref["anc_hist"][t][diff])
            assert shift.max() == 1, (d, n, t, "not a boundary tie", shift.max())
            assert ties <= max(1, n // 100000), (d, n, t, ties)
            break
    print(f"lgmvn d={d} n={n}: boundary ties {ties} (first at step {tie_step}) of {n * T} pointers")
    if ties == 0:
        assert np.array_equal(anc, ref["anc_hist"])
        assert np.allclose(xh[T], ref["x_hist"][T], rtol=1e-9, atol=1e-9)
        assert np.allclose(out.x_final, ref["x_final"], rtol=1e-9, atol=1e-9)      # every particle


def test_lgmvn_log_marginal_vs_kalman():
    """log-marginal vs the exact Kalman filter.  A bootstrap filter degenerates
    with the dimension (at d = 64 its estimate is far in the lower tail with any
    feasible N), so the exact check runs at d = 8; the d = 64 shape of BASELINE
    config 5 is pinned number for number against the oracle sweep above."""
    from anglican_b200 import Population
    from anglican_b200.models import LinearGaussianMVN
    from oracle import pyoracle as orc
    d, T, n = 8, 8, 1 << 18
    A, Qc, R, y = _lgmvn(d, T)
    m = LinearGaussianMVN(A, Qc, R)
    om = orc.Model(orc.SSM_LGMVN, dim=d, s0=1.0, A=A, Lq=np.linalg.cholesky(Qc), R=R)
    with Population(n, seed=5) as pop:
        out = pop.sweep(m, y.ravel(), want_final=False)
    exact, _ = orc.kalman_mvn(om, y.ravel())
    assert abs(out.log_z - exact) < 0.5, (out.log_z, exact)

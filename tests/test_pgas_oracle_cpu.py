"""Particle Gibbs with ancestor sampling (pgas.cljc:107-168) in the oracle
(`ancestor_sampling` of orc_smc_sweep): with very few particles and a long
series, plain particle Gibbs never moves the early states of the retained path
(path degeneracy), ancestor sampling does and reproduces the exact smoothing
distribution (Rauch-Tung-Striebel smoother of the linear-Gaussian model).
The CUDA path is compared with this oracle in tests/test_zz_pgas_gpu.py."""
import numpy as np

from oracle import pyoracle as orc

PAR = dict(m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)


def rts_smoother(y, m0, s0, a, b, q, c, d, r):
    T = len(y)
    mf, Pf, mp, Pp = np.zeros(T + 1), np.zeros(T + 1), np.zeros(T + 1), np.zeros(T + 1)
    mf[0], Pf[0] = m0, s0 * s0
    for t in range(1, T + 1):
        mp[t], Pp[t] = a * mf[t - 1] + b, a * a * Pf[t - 1] + q * q
        K = Pp[t] * c / (c * c * Pp[t] + r * r)
        mf[t] = mp[t] + K * (y[t - 1] - c * mp[t] - d)
        Pf[t] = (1 - K * c) * Pp[t]
    ms, Ps = mf.copy(), Pf.copy()
    for t in range(T - 1, -1, -1):
        G = Pf[t] * a / Pp[t + 1]
        ms[t] = mf[t] + G * (ms[t + 1] - mp[t + 1])
        Ps[t] = Pf[t] + G * G * (Ps[t + 1] - Pp[t + 1])
    return ms, Ps


def lgssm_series(T, seed=5):
    rng = np.random.default_rng(seed)
    x, ys = 0.0, []
    for _ in range(T):
        x = PAR["a"] * x + rng.normal()
        ys.append(x + rng.normal())
    return np.array(ys)


def run_chain(y, n, iters, burn, ancestor_sampling, seed0=1000):
    om = orc.Model(orc.SSM_LGSSM1D, **PAR)
    T = y.size
    g = np.random.default_rng(1)
    retained, prev = None, None
    acc, acc2, moved = np.zeros(T + 1), np.zeros(T + 1), np.zeros(T + 1)
    for it in range(iters):
        cond = retained is not None
        r = orc.smc_sweep(om, y, n, seed=seed0 + it, retained=retained, csmc=cond, n_cond_steps=T,
                          reset_weights=True, keep_history=True,
                          ancestor_sampling=ancestor_sampling and cond)
        retained = orc.trace_lineage(r, int(g.random() * n))[:, 0].copy()
        if it >= burn:
            acc += retained
            acc2 += retained ** 2
            if prev is not None:
                moved += retained != prev
        prev = retained
    k = iters - burn
    mean = acc / k
    return mean, acc2 / k - mean * mean, moved / (k - 1)


def test_ancestor_sampling_mixes_and_matches_the_smoother():
    y = lgssm_series(40)
    ms, Ps = rts_smoother(y, **PAR)
    mean, var, moved = run_chain(y, 6, 4000, 200, True)
    assert np.abs(mean - ms).max() < 0.15
    assert np.all(np.abs(var / Ps - 1.0) < 0.2)
    assert moved.min() > 0.1 and moved.mean() > 0.5     # every x_t keeps being refreshed
    # plain particle Gibbs with the same budget is stuck on the early states
    _, _, moved_pg = run_chain(y, 6, 600, 100, False)
    assert moved_pg[:10].max() < 0.05 and moved_pg[-1] > 0.4


def test_ancestor_draw_only_changes_the_retained_ancestry():
    """forward particles, weights and log Z are those of the conditional sweep"""
    om = orc.Model(orc.SSM_LGSSM1D, **PAR)
    y = lgssm_series(15)
    first = orc.smc_sweep(om, y, 50, seed=7, keep_history=True, reset_weights=True)
    ret = orc.trace_lineage(first, 3)[:, 0].copy()
    kw = dict(retained=ret, csmc=True, n_cond_steps=15, reset_weights=True, keep_history=True)
    a = orc.smc_sweep(om, y, 50, seed=8, **kw)
    b = orc.smc_sweep(om, y, 50, seed=8, ancestor_sampling=True, **kw)
    assert a["log_z"] == b["log_z"] and np.array_equal(a["x_hist"], b["x_hist"])
    assert np.array_equal(a["anc_hist"][:, 1:], b["anc_hist"][:, 1:])
    assert np.all(a["anc_hist"][:, 0] == 0) and np.any(b["anc_hist"][:, 0] != 0)
    # HMM: transition rows as the rescored density
    hm = orc.Model(orc.SSM_HMM, dim=3, hmm_init=[1.0, 1.0, 1.0],
                   hmm_trans=[[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]],
                   hmm_mu=[-1.0, 1.0, 0.0], hmm_sd=[1.0, 1.0, 1.0])
    yh = np.array([0.9, 0.8, 0.7, 0.0, -0.025, -5.0, -2.0, -0.1])
    f = orc.smc_sweep(hm, yh, 40, seed=3, keep_history=True, reset_weights=True)
    rt = orc.trace_lineage(f, 0)[:, 0].copy()
    h = orc.smc_sweep(hm, yh, 40, seed=4, retained=rt, csmc=True, n_cond_steps=8,
                      reset_weights=True, keep_history=True, ancestor_sampling=True)
    assert np.any(h["anc_hist"][:, 0] != 0) and np.all(h["anc_hist"][:, 0] < 40)

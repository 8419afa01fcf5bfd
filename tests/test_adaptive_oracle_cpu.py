"""The oracle's adaptive (ESS-threshold) resampling rule - an engine extension, the
reference resamples at every observe (smc.cljc:41-43) - pinned on its two limits and on
an exact truth, so that the GPU parity test (tests/test_zz_adaptive_gpu.py) compares
against something that is itself checked:
  threshold -> 0+   never resamples      == `infer :importance` (importance.cljc:16-22)
  threshold  = 1    resamples whenever ESS < N, i.e. (almost) always == the reference rule
  in between        log Z-hat still estimates the Kalman log-marginal."""
import numpy as np

from oracle import pyoracle as orc


def _lgssm(T, seed=20240601):
    rng = np.random.default_rng(seed)
    x, ys = rng.normal(), []
    for _ in range(T):
        x = 0.9 * x + rng.normal()
        ys.append(x + rng.normal())
    return np.array(ys)


M = orc.Model(orc.SSM_LGSSM1D, m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)


def test_threshold_near_zero_is_importance_sampling():
    y = _lgssm(12)
    a = orc.smc_sweep(M, y, 3000, seed=4, ess_threshold=1e-12)
    b = orc.importance_sweep(M, y, 3000, seed=4)
    assert np.all(a["log_mean_w"] == 0.0)                        # no observe resampled
    assert np.array_equal(a["x_final"][:, 0], b["x_final"])
    assert np.allclose(a["lw_final"], b["lw_final"], rtol=0, atol=1e-12)
    assert abs(a["log_z"] - b["log_z"]) < 1e-12


def test_threshold_one_is_the_reference_rule():
    y = _lgssm(15)
    a = orc.smc_sweep(M, y, 2000, seed=9, ess_threshold=1.0, keep_history=True)
    b = orc.smc_sweep(M, y, 2000, seed=9, keep_history=True)
    assert a["log_z"] == b["log_z"] and np.array_equal(a["anc_hist"], b["anc_hist"])
    assert np.array_equal(a["x_final"], b["x_final"])
    assert np.all(a["lw_final"] == a["log_z"])                   # smc.cljc:98-103: one common weight


def test_adaptive_log_z_estimates_the_kalman_marginal():
    y = _lgssm(40)
    exact = orc.kalman1d(M, y)[0]
    for thr in (0.3, 0.7):
        zs = [orc.smc_sweep(M, y, 20000, seed=s, ess_threshold=thr, threads=8)["log_z"] for s in range(3)]
        assert abs(np.mean(zs) - exact) < 0.15, (thr, zs, exact)
    r = orc.smc_sweep(M, y, 20000, seed=1, ess_threshold=0.5)
    kept = int((r["log_mean_w"] == 0.0).sum())
    assert 0 < kept < y.size                                     # both branches taken
    # deferred increments: the sum over the resampling observes (+ the final term) is log Z
    w = r["lw_final"]
    if np.ptp(w) == 0.0:
        assert abs(r["log_mean_w"].sum() - r["log_z"]) < 1e-9

"""Parity of the CUDA resampler (through the C ABI) with the oracle's literal
restatement of smc.cljc:86-137.  Bar: ancestor indices bit-exact for the same
fp64 weights and uniform; mismatches are reported with their distance to the
nearest cumulative edge (boundary ties) and fail the test."""
import math

import numpy as np
import pytest

from tests.weights_cases import U_VALUES, families

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orc():
    from oracle import pyoracle
    return pyoracle


def _pop(n, seed=1):
    from anglican_b200 import Population
    return Population(n, seed=seed)


def _report(w, u, n_new, got_emission, ref):
    bad = np.nonzero(got_emission != ref["emission"])[0]
    edges = np.cumsum(w)
    step = ref["step"]
    x = ref["total"] * u + step * bad[:10]
    d = [float(np.min(np.abs(edges - (xx if xx < ref["total"] else xx - ref["total"]))))
         for xx in x]
    return f"{bad.size} mismatches; first at {bad[:10]}; |x-edge|/total={np.array(d)/ref['total']}"


@pytest.mark.parametrize("n", [1, 2, 3, 17, 255, 256, 257, 4095, 4096, 4097, 8192, 20000, 100003])
def test_systematic_weights_bit_exact(orc, n):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(1000 + n)
    with _pop(n) as pop:
        wcol = pop.column(L.F64)
        anc = pop.column(L.I32)
        for name, w in families(n, rng).items():
            w = np.ascontiguousarray(w, dtype=np.float64)
            wcol.upload(w)
            for u in U_VALUES:
                for n_new in sorted({n, max(1, n - 1), max(1, n // 3)}):
                    ref = orc.resample_weights(w, n_new, u)
                    _, total, J = pop.resample_weights(wcol, u=u, n_new=n_new,
                                                       order=L.ORDER_EMISSION, anc=anc)
                    got = anc.download()[:n_new]
                    assert total == ref["total"], (name, u, n_new, total, ref["total"])
                    assert J == int(np.sum(ref["lap"] == 0)), (name, u, n_new)
                    assert np.array_equal(got, ref["emission"]), \
                        (name, u, n_new, _report(w, u, n_new, got, ref))


@pytest.mark.parametrize("n", [5, 1000, 4097, 50000])
def test_orders(orc, n):
    """REFERENCE order == the list smc.cljc returns; SORTED == engine physical
    order (second-lap emissions first), both closed-form permutations."""
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(77 + n)
    with _pop(n) as pop:
        lw = rng.normal(0, 2.5, n)
        w = np.exp(lw - lw.max())
        wcol = pop.column(L.F64, data=w)
        for u in [0.0, 0.41, 0.93]:
            for n_new in [n, n - 1]:
                ref = orc.resample_weights(w, n_new, u)
                a_ref, _, J = pop.resample_weights(wcol, u=u, n_new=n_new, order=L.ORDER_REFERENCE)
                assert np.array_equal(a_ref.download()[:n_new], ref["anc_ref"])
                a_s, _, _ = pop.resample_weights(wcol, u=u, n_new=n_new, order=L.ORDER_SORTED)
                srt, j0 = orc.emission_to_sorted(ref["emission"], ref["lap"])
                got = a_s.download()[:n_new]
                assert J == j0
                assert np.array_equal(got, srt)
                assert np.all(np.diff(got) >= 0), "sorted order must be monotone"


@pytest.mark.parametrize("n", [3, 1000, 65536, 300001])
def test_resample_from_logweights(orc, n):
    """apop_resample: particle-weights (smc.cljc:70-84) + wheel + new log-weight
    (smc.cljc:98-103)."""
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(5 + n)
    with _pop(n) as pop:
        for kind in ["normal", "with_neg_inf", "all_neg_inf", "offset", "with_pos_inf"]:
            lw = rng.normal(-3.0, 2.0, n)
            if kind == "with_pos_inf":
                # max = +inf: weight 1 for the particles AT the max, 0 for all others, and the
                # common log-weight afterwards is -inf "so that the sweep dies out"
                # (smc.cljc:78-82, 98-103)
                lw[rng.integers(0, n, size=max(1, n // 50))] = np.inf
                lw[rng.random(n) < 0.1] = -np.inf
                lw[n // 2] = np.inf
            elif kind == "with_neg_inf":
                lw[rng.random(n) < 0.3] = -np.inf
                lw[0] = -1.0
            elif kind == "all_neg_inf":
                lw[:] = -np.inf
            elif kind == "offset":
                lw += -1234.5
            col = pop.column(L.F64, data=lw)
            for u in [0.0, 0.6180339887]:
                ref_anc, ref_lw = orc.resample(lw, n, u)
                anc, new_lw, _ = pop.resample(col, u=u, order=L.ORDER_REFERENCE)
                got = anc.download()
                nbad = int(np.sum(got != ref_anc))
                assert nbad == 0, (kind, u, nbad)
                if math.isinf(ref_lw):
                    assert new_lw == ref_lw
                else:
                    assert abs(new_lw - ref_lw) <= 1e-12 * max(1.0, abs(ref_lw))
            col.free()


def test_nan_and_zero_weights_are_errors():
    from anglican_b200 import _lib as L
    with _pop(1000) as pop:
        lw = np.zeros(1000)
        lw[17] = np.nan
        col = pop.column(L.F64, data=lw)
        with pytest.raises(L.ApopError) as e:
            pop.resample(col, u=0.5)
        assert e.value.code == -3          # reference: resample never terminates
        w = pop.column(L.F64, data=np.zeros(1000))
        with pytest.raises(L.ApopError) as e:
            pop.resample_weights(w, u=0.5)
        assert e.value.code == -4


def test_normalize_matches_oracle(orc):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(3)
    n = 123457
    lw = rng.normal(-50, 4, n)
    with _pop(n) as pop:
        col = pop.column(L.F64, data=lw)
        m, lm, ess = pop.normalize(col)
        assert m == lw.max()
        assert abs(lm - orc.log_marginal(lw)) <= 1e-9 * abs(lm)
        assert abs(ess - orc.ess(lw)) <= 1e-9 * ess


def test_multinomial_supplied_uniforms(orc):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(11)
    n = 70001
    w = rng.random(n) ** 3
    u = rng.random(n)
    with _pop(n) as pop:
        wc = pop.column(L.F64, data=w)
        uc = pop.column(L.F64, data=u)
        anc = pop.resample_multinomial_u(wc, uc)
        assert np.array_equal(anc.download(), orc.multinomial_weights(w, u))


def test_gather(orc):
    from anglican_b200 import _lib as L
    rng = np.random.default_rng(12)
    n = 10007
    x = rng.normal(size=(n, 3))
    idx = rng.integers(0, n, n).astype(np.int32)
    with _pop(n) as pop:
        xc = pop.column(L.F64, width=3, data=x)
        ic = pop.column(L.I32, data=idx)
        pop.gather(ic, [xc])
        assert np.array_equal(xc.download(), x[idx])


@pytest.mark.parametrize("logn", [24, 27])
def test_large_population_bit_exact(orc, logn):
    """BASELINE sizes (C3 shard sizes 2^24 and the whole 2^27 population): the
    oracle's sequential wheel still finishes in seconds."""
    from anglican_b200 import _lib as L
    n = 1 << logn
    rng = np.random.default_rng(7)          # SURVEY §8(d) microbench: lw ~ N(0, 3^2), rng(7)
    lw = rng.normal(0, 3, n)
    w = np.exp(lw - lw.max())
    u = 0.3141592653589793
    with _pop(n) as pop:
        wc = pop.column(L.F64, data=w)
        anc, total, J = pop.resample_weights(wc, u=u, order=L.ORDER_EMISSION)
        got = anc.download()
    ref = orc.resample_weights(w, n, u)
    assert total == ref["total"]
    nbad = int(np.sum(got != ref["emission"]))
    assert nbad == 0, f"{nbad} ancestor mismatches at n=2^{logn}"


@pytest.mark.parametrize("case", ["one_giant", "forty_heavy", "giant_plus_background", "logweights"])
def test_degenerate_weights_deferred_runs(orc, case):
    """Weights so peaked that one tile owns more than HEAVY_W (2^18) offspring:
    k_ancestors hands those runs to k_fill_runs instead of emitting them from
    one CTA.  Ancestors stay bit-exact in every output order."""
    from anglican_b200 import _lib as L
    n = 1 << 20
    rng = np.random.default_rng(5)
    w = np.full(n, 1e-10)
    if case == "one_giant":
        w[777777] = 1.0                                   # one run of ~2^20 (>= BIG_RUN)
    elif case == "forty_heavy":
        w[300000:300040] = 1.0 + rng.random(40)           # 40 runs of ~26000 in one tile
    elif case == "giant_plus_background":
        w = rng.random(n)
        w[123] = 3.0 * n                                  # 3/4 of the offspring + ordinary tiles
    else:
        w = None
    with _pop(n) as pop:
        anc = pop.column(L.I32)
        if w is not None:
            wcol = pop.column(L.F64, data=w)
            for u in (0.0, 0.37, 0.999999):
                ref = orc.resample_weights(w, n, u)
                for order, key in ((L.ORDER_EMISSION, "emission"), (L.ORDER_REFERENCE, "anc_ref")):
                    _, total, _ = pop.resample_weights(wcol, u=u, n_new=n, order=order, anc=anc)
                    assert total == ref["total"]
                    assert np.array_equal(anc.download(), ref[key]), (case, u, order)
        else:
            lw = rng.normal(0, 1, n)
            lw[999] = 40.0                                # exp(40) against ~1: one giant particle
            lcol = pop.column(L.F64, data=lw)
            ref_anc, ref_lw = orc.resample(lw, n, 0.42)      # reference order (smc.cljc:127)
            _, new_lw, _ = pop.resample(lcol, u=0.42, order=L.ORDER_REFERENCE, anc=anc)
            assert np.array_equal(anc.download(), ref_anc)
            assert abs(new_lw - ref_lw) <= 1e-9 * abs(ref_lw)


@pytest.mark.parametrize("n", [4097, 20000, 100003])
def test_deferred_runs_forced_at_small_sizes(orc, n, monkeypatch):
    """APOP_HEAVY_W lowers the threshold so that ordinary tiles take the deferred
    (run list + k_fill_runs) path: every weight family, every order, bit-exact"""
    from anglican_b200 import _lib as L
    monkeypatch.setenv("APOP_HEAVY_W", "1500")
    rng = np.random.default_rng(n)
    with _pop(n) as pop:
        wcol = pop.column(L.F64)
        anc = pop.column(L.I32)
        for name, w in families(n, rng).items():
            w = np.ascontiguousarray(w, dtype=np.float64)
            wcol.upload(w)
            for u in (0.0, 0.6180339887):
                ref = orc.resample_weights(w, n, u)
                srt = orc.emission_to_sorted(ref["emission"], ref["lap"])
                ref["anc_sorted"] = srt[0] if isinstance(srt, tuple) else srt
                for order, key in ((L.ORDER_EMISSION, "emission"), (L.ORDER_REFERENCE, "anc_ref"),
                                   (L.ORDER_SORTED, "anc_sorted")):
                    pop.resample_weights(wcol, u=u, n_new=n, order=order, anc=anc)
                    assert np.array_equal(anc.download(), ref[key]), (name, u, key)

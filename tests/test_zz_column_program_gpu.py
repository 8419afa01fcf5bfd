"""The general lock-step path, end to end (DESIGN §1 "column level"):
BASELINE config 1 - the HMM of algorithm_test.cljc:116-147 - run as a PROGRAM OVER
COLUMNS, one C-ABI call per `sample*` / `observe*` / `resample` / gather of the
traced program (inference.cljc:42-50, smc.cljc:35-51, 86-137), against the same
program executed particle by particle from the oracle's primitives:

    (let [s0 (sample* (discrete init))]                       apop_sample_discrete
      (loop [s s0, ys obs]
        (let [s' (sample* (discrete (nth trans s)))]           apop_column_take + apop_sample_discrete
          (observe* (normal (nth mu s') 1) (first ys))        apop_column_take + apop_observe_normal
          (recur s' (rest ys)))))                             apop_resample + apop_gather at the observe

States and ancestors are integers: bit-exact.  log Z-hat: 1e-9 relative."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

OBS = [0.9, 0.8, 0.7, 0.0, -0.025, -5.0, -2.0, -0.1, 0.0, 0.13, 0.45, 6, 0.2, 0.3, -1, -1]
INIT = [1.0, 1.0, 1.0]
TRANS = [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]]
MU = [-1.0, 1.0, 0.0]
STREAM_USER, STREAM_RESAMPLE = 3, 2


def _discrete(w, u):
    """runtime.cljc:277-287"""
    total = 0.0
    for v in w:
        total = total + v
    x, acc = u * total, 0.0
    for i, v in enumerate(w):
        acc = acc + v
        if x < acc:
            return i
    return len(w) - 1


def _reference_program(orc, n, seed):
    """the sweep of smc.cljc:35-51 over n particles, one particle at a time"""
    s = np.array([_discrete(INIT, orc.uniform_at(seed, k, 0, STREAM_USER, 0)) for k in range(n)])
    lw = np.zeros(n)
    ancs = []
    for t, y in enumerate(OBS, start=1):
        s = np.array([_discrete(TRANS[s[k]], orc.uniform_at(seed, k, t, STREAM_USER, 0))
                      for k in range(n)])
        lw = lw + np.array([orc.normal_logpdf(y, MU[s[k]], 1.0) for k in range(n)])
        u = orc.uniform_at(seed, 0, t, STREAM_RESAMPLE, 0)
        anc, new_lw = orc.resample(lw, n, u)                  # reference order (smc.cljc:127)
        ancs.append(anc)
        s = s[anc]
        lw = np.full(n, new_lw)
    return s, float(lw[0]), ancs


@pytest.mark.parametrize("n", [1000, 5000])
def test_hmm_as_a_program_over_columns(n):
    from anglican_b200 import Population, _lib as L
    from oracle import pyoracle as orc
    seed = 4242
    with Population(n, seed=seed) as pop:
        s = pop.column(L.I32)
        rows = pop.column(L.F64, width=3)
        mean = pop.column(L.F64)
        lw = pop.column(L.F64, fill=0.0)
        anc = pop.column(L.I32)
        pop.set_step(0)
        pop.sample_discrete(INIT, s, draw=0)
        got_anc = []
        for t, y in enumerate(OBS, start=1):
            pop.set_step(t)
            pop.take(TRANS, s, rows)                           # (nth trans s)
            pop.sample_discrete(rows, s, draw=0)
            pop.take(MU, s, mean)                              # (nth mu s')
            pop.observe_normal(mean, 1.0, float(y), lw)
            _, new_lw, _ = pop.resample(lw, order=L.ORDER_REFERENCE, anc=anc)
            got_anc.append(anc.download())
            pop.gather(anc, [s])
            lw.fill(new_lw)                                    # smc.cljc:98-103, 128
        s_final = s.download()
        log_z = new_lw
    ref_s, ref_logz, ref_anc = _reference_program(orc, n, seed)
    for t, (a, b) in enumerate(zip(got_anc, ref_anc)):
        assert np.array_equal(a, b), ("ancestors", t)
    assert np.array_equal(s_final, ref_s)
    assert abs(log_z - ref_logz) <= 1e-9 * abs(ref_logz)
    # and the program estimates the right thing: exact forward log Z = -44.42446961347636
    if n >= 5000:
        assert abs(log_z - (-44.42446961347636)) < 0.5


def test_gather_widths_and_types():
    """apop_gather: width-1 fast path, 16-byte rows, odd rows, i32; n_new < n keeps the tail"""
    from anglican_b200 import Population, _lib as L
    n = 10007
    rng = np.random.default_rng(3)
    idx = np.sort(rng.integers(0, n, n)).astype(np.int32)
    with Population(n, seed=1) as pop:
        anc = pop.column(L.I32, data=idx)
        for dtype, width in [(L.F64, 1), (L.I32, 1), (L.F64, 2), (L.F64, 3), (L.F64, 8), (L.I32, 4),
                             (L.I32, 5)]:
            npd = np.float64 if dtype == L.F64 else np.int32
            data = rng.integers(0, 1 << 20, (n, width)).astype(npd)
            col = pop.column(dtype, width=width, data=data)
            pop.gather(anc, [col])
            got = col.download().reshape(n, width)
            assert np.array_equal(got, data[idx]), (dtype, width)
            col2 = pop.column(dtype, width=width, data=data)
            pop.gather(anc, [col2], n_new=n - 1)               # conditional SMC: N-1 offspring
            got = col2.download().reshape(n, width)
            assert np.array_equal(got[:n - 1], data[idx[:n - 1]]) and np.array_equal(got[n - 1], data[n - 1])

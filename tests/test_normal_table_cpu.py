"""The Gaussian draw of the engine: 64 random bits -> z through a table-driven
inverse normal CDF (csrc/rng.cuh `normal_core`/`normal_tail`, restated in
oracle/anglican_oracle.c `orc_normal_from_bits`).  The reference's nextGaussian
(runtime.cljc:403-409 on the unseeded RNG of 142-147) is only distributionally
comparable, so the map is checked against the exact quantile function."""
import os

import numpy as np
from scipy.special import ndtri
from scipy import stats

from oracle import pyoracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _v_of(hi, lo):
    m = ((int(hi) & 0xFFFFF) << 32) | int(lo)
    return m / 2.0 ** 53, ((int(hi) >> 20) & 0x7FF)


def test_engine_and_oracle_tables_are_the_same_file():
    a = open(os.path.join(ROOT, "anglican_b200", "csrc", "normal_table.inc")).read()
    b = open(os.path.join(ROOT, "oracle", "normal_table.inc")).read()
    assert a == b


def test_quantile_map_matches_ndtri_everywhere():
    rng = np.random.default_rng(11)
    worst = 0.0
    his = rng.integers(0, 2 ** 32, 200000, dtype=np.uint64)
    los = rng.integers(0, 2 ** 32, 200000, dtype=np.uint64)
    # force every octave (and the tail) to be visited: clear leading mantissa bits
    for k in range(200000):
        sh = k % 40
        if sh:
            m = (((int(his[k]) & 0xFFFFF) << 32) | int(los[k])) >> sh
            his[k] = (int(his[k]) & 0xFFF00000) | (m >> 32)
            los[k] = m & 0xFFFFFFFF
    for hi, lo in zip(his, los):
        z = orc.normal_from_bits(hi, lo)
        v, ext = _v_of(hi, lo)
        tail = v < 2.0 ** -17
        ve = v + (ext + 0.5) * 2.0 ** -64 if tail else v
        if ve == 0.0:
            continue
        ref = -ndtri(ve)
        sign = -1.0 if (int(hi) >> 31) else 1.0
        err = abs(z - sign * ref)
        tol = 1e-11 if not tail else 1e-9 * abs(ref)
        assert err <= tol, (hex(int(hi)), hex(int(lo)), z, sign * ref)
        worst = max(worst, err)
    assert worst < 1e-11 or worst < 1e-8


def test_rows_join_continuously():
    # v just below / at a row boundary: the two polynomials agree to the fit error
    for e in range(1, 17):
        for i in (0, 1, 17, 31):
            v = 2.0 ** -(e + 1) * (1 + i / 32)
            m = int(v * 2 ** 53)
            a = orc.normal_from_bits(m >> 32, m & 0xFFFFFFFF)
            m2 = m - 1
            b = orc.normal_from_bits(m2 >> 32, m2 & 0xFFFFFFFF)
            if v - 2.0 ** -53 >= 2.0 ** -17:
                assert abs(a - b) < 2e-11 + 2.0 ** -53 / stats.norm.pdf(a) * 4, (e, i, a, b)


def test_draws_are_standard_normal():
    n = 400000
    z = np.array([orc.normal_at(0xA11CE, k, 3, 1, 0) for k in range(n)])
    assert abs(z.mean()) < 4 / np.sqrt(n)
    assert abs(z.var() - 1.0) < 4 * np.sqrt(2.0 / n)
    assert abs(stats.skew(z)) < 4 * np.sqrt(6.0 / n)
    assert abs(stats.kurtosis(z)) < 4 * np.sqrt(24.0 / n)
    assert stats.kstest(z, "norm").pvalue > 1e-3
    # slots of a pair are different draws
    assert abs(np.corrcoef(z[0::2], z[1::2])[0, 1]) < 4 / np.sqrt(n / 2)

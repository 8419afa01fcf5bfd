"""One particle outweighing every other by more than 2^53 (each other increment
vanishes against the running sum), anywhere in the population.  The reference's
wheel (smc.cljc:86-137) then emits copies of that particle only (plus whatever
lies below the first pointer).  Regression for the tile classification: the
approximate exclusive prefix must stay relatively accurate next to a dominant
tile sum (exclusive values taken from the neighbour's inclusive value, never by
subtracting the own value), otherwise easy tiles before the heavy one were put
into the wrong binade and the chain's consistency check fired
(APOP_ERR_INTERNAL bits 0x204 / 0x604).  Shapes from tools/dbg_giant.py.

Named test_zz_* so that it runs after the rest of the GPU suite."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TILE = 4096

CASES = [
    # n, background, position of the heavy particle, its weight, random background
    (450000, 1e-30, 270000, 1.0, False),
    (600000, 1e-30, 360000, 1.0, False),     # tile 87 = 4*21 + 3
    (600000, 1e-30, 570000, 1.0, False),     # tile 139: beyond the first warp of the tile scan
    (800000, 1e-30, 760000, 1.0, False),
    (1 << 20, 1e-30, 629145, 1.0, False),
    (1 << 20, 1e-30, 996147, 1.0, False),
    (1 << 20, 1e-30, 777777, 1.0, True),
    (1 << 20, 1e-30, 777777, 1.5, False),    # not a power of two: tiles after it stay easy
    (1 << 20, 1e-20, 777777, 1.5, True),
    (1 << 20, 1e-5, 777777, 1e6, False),     # ratio 1e11: prefix survives in part
    (1 << 20, 1e-9, 777777, 1e6, True),      # ratio 1e15: prefix nearly cancelled
    (3 * TILE, 1e-30, 3 * TILE - 1, 1.0, False),
    (1 << 22, 1e-30, (1 << 22) - 5, 3.0, False),
]


@pytest.mark.parametrize("n,small,pos,giant,rnd", CASES)
@pytest.mark.parametrize("u", [0.37, 0.999999])
def test_dominant_weight_any_tile(n, small, pos, giant, rnd, u):
    from anglican_b200 import Population, _lib as L
    from oracle import pyoracle as orc
    rng = np.random.default_rng(n + pos)
    w = np.full(n, small)
    if rnd:
        w *= rng.random(n) + 0.5
    w[pos] = giant
    ref = orc.resample_weights(w, n, u)
    with Population(n) as pop:
        wcol = pop.column(L.F64, data=w)
        anc = pop.column(L.I32)
        _, total, _ = pop.resample_weights(wcol, u=u, n_new=n, order=L.ORDER_EMISSION, anc=anc)
        got = anc.download()
    assert total == ref["total"]
    assert np.array_equal(got, ref["emission"])
    if small * 1.5 * n < giant * 2.0 ** -53:        # the rest never moves the running sum
        assert np.all(got == pos)
